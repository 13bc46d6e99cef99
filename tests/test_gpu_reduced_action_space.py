"""reduced_action_space (reference shadow_hand.py:73-79,160-181; shadow_hand_test.py): the joints THJ5, THJ1, LFJ5 of
each hand and their actuators leave the model - here they are locked (scene/compile.py) and dropped from the specs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_reduced_action_space_specs_and_parity():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import PianoWithShadowHandsOracle

    midi = music.load("TwinkleTwinkleRousseau")
    task = suite.PianoWithShadowHands(midi, n_steps_lookahead=1, trim_silence=True, reduced_action_space=True)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=True, auto_reset=False)
    # reference shadow_hand_test.py: 3 actuators and 3 joints fewer per hand
    assert env.action_spec().shape == (45 - 6,)
    spec = env.observation_spec()
    assert spec["rh_shadow_hand/joints_pos"].shape == (23,) and spec["lh_shadow_hand/joints_pos"].shape == (23,)
    names = task.model.names
    assert [names["dof"][d].split("_")[-1] for d in task.model.locked_dofs] == ["LFJ5", "THJ5", "THJ1"] * 2
    thj2 = next(i for i, n in enumerate(names["dof"]) if n.endswith("rh_THJ2"))
    np.testing.assert_allclose(task.model.a["jnt_range"][thj2], (0.0, 0.698132))

    ora = PianoWithShadowHandsOracle(task)  # same locked tables, full 45-wide action
    act_keep, obs_keep, _ = task.reduced_spec()
    obs = env.reset()
    ora.reset()
    rng = np.random.RandomState(0)
    qpos = torch.zeros((2, 140), device="cuda")
    for step in range(6):
        a = rng.uniform(-1, 1, 39)
        full = np.zeros(45)
        full[act_keep] = a
        obs, rew, done = env.step(np.stack([a, a]))
        obs_o, r_o, _ = ora.step(ora.canonical_to_spec(full.astype(np.float32)))
        for h in ("rh", "lh"):
            k = f"{h}_shadow_hand/joints_pos"
            joints_o = obs_o[k]
            lanes = [d - 88 - (0 if h == "rh" else 26) for d in task.model.locked_dofs]
            drop = {(l - 2 if l >= 2 else 24 + l) for l in lanes if 0 <= l < 26}
            keep = [j for j in range(26) if j not in drop]
            np.testing.assert_allclose(obs[k][0], joints_o[keep], atol=3e-4)
        np.testing.assert_allclose(rew[0], r_o, atol=5e-3)
        env.engine.get_state(qpos=qpos)
        torch.cuda.synchronize()
        assert qpos[:, task.model.locked_dofs].abs().max().item() < 1e-6  # the locked joints do not move
    env.close()

    # the full hand under the same 39 commands (dead actuators at mid range) moves differently
    task_full = suite.PianoWithShadowHands(midi, n_steps_lookahead=1, trim_silence=True)
    env_full = suite.BatchedEnvironment(task_full, 1, canonical_actions=True, auto_reset=False)
    env_full.reset()
    rng = np.random.RandomState(0)
    for step in range(6):
        a = rng.uniform(-1, 1, 39)
        full = np.zeros(45)
        full[act_keep] = a
        obs_f, _, _ = env_full.step(full[None])
    assert np.abs(obs_f["rh_shadow_hand/joints_pos"][0][[j for j in range(26) if j not in drop]]
                  - obs["rh_shadow_hand/joints_pos"][0]).max() > 1e-3
    env_full.close()


def test_reduced_device_api_matches_numpy_api():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diffusion_piano_b200 import music, suite

    task = suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), n_steps_lookahead=1, reduced_action_space=True)
    e1 = suite.BatchedEnvironment(task, 3, canonical_actions=True)
    e2 = suite.BatchedEnvironment(task, 3, canonical_actions=True)
    assert e1.obs_dim == e1.engine.obs_dim - 6 and e1.action_dim == 39
    o1 = e1.reset_device().cpu().numpy().copy()
    o2 = e2.reset()
    rng = np.random.RandomState(1)
    for _ in range(3):
        a = rng.uniform(-1, 1, (3, 39)).astype(np.float32)
        o1, r1, d1, _, _ = e1.step_device(torch.as_tensor(a, device="cuda"))
        o2, r2, d2 = e2.step(a)
        flat2 = np.concatenate([o2[k] for k, _ in sorted(e2.layout.items(), key=lambda kv: kv[1][0])], axis=1)
        np.testing.assert_array_equal(o1.cpu().numpy().astype(np.float64), flat2)
        np.testing.assert_array_equal(r1.cpu().numpy().astype(np.float64), r2)
    e1.close()
    e2.close()
