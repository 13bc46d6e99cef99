"""PianoWithOneShadowHand (reference robopianist/suite/tasks/piano_with_one_shadow_hand.py:40-388) on the GPU:
the other hand is locked, parked and dead inside the two-hand step kernel (tests/test_scene_and_oracle.py shows
that form equal to the reference's one-hand model); here the CUDA step + facade against the fp64 oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("side,fingering", [("RIGHT", True), ("LEFT", True), ("RIGHT", False)])
def test_one_hand_specs_and_parity(side, fingering):
    import torch
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import PianoWithShadowHandsOracle

    midi = music.load("CMajorScaleTwoHands")  # both hands' notes: the one hand is asked to play all of them
    hs = getattr(suite.HandSide, side)
    pre = "lh" if side == "LEFT" else "rh"
    task = suite.PianoWithOneShadowHand(midi, hs, n_steps_lookahead=2, disable_fingering_reward=not fingering)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=False, auto_reset=False)
    # action: the hand's 22 actuators + sustain (:190-200); observables (:320-361)
    assert env.action_spec().shape == (23,)
    spec = env.observation_spec()
    want = {"goal": 3 * 89, "piano/state": 88, "piano/sustain_state": 1, f"{pre}_shadow_hand/joints_pos": 26,
            f"{pre}_shadow_hand/position": 3}
    if fingering:
        want["fingering"] = 5
    assert {k: v.shape[0] for k, v in spec.items()} == want
    assert task.reward_names == ("key_press_reward", "sustain_reward", "energy_reward") + \
        (("fingering_reward",) if fingering else ())
    ora = PianoWithShadowHandsOracle(task)
    act_keep = task.reduced_spec()[0]
    obs, obs_o = env.reset(), ora.reset()
    for k, v in obs_o.items():
        np.testing.assert_allclose(obs[k][0], v, atol=1e-6, err_msg=k)
    rng = np.random.RandomState(1)
    lo, hi = env.action_spec().minimum, env.action_spec().maximum
    cnt = torch.zeros((2, 8), dtype=torch.int32, device="cuda")
    seen_fingering = set()
    for step in range(16):
        a = lo + rng.uniform(0.25, 0.75, 23) * (hi - lo)
        full = np.zeros(45)
        full[act_keep] = a
        obs, rew, done = env.step(np.stack([a, a]))
        obs_o, r_o, d_o = ora.step(np.asarray(full, np.float32).astype(np.float64))
        env.engine.get_diag(counters=cnt)
        torch.cuda.synchronize()
        assert int(cnt[0, 3]) == 0
        assert set(obs) == set(obs_o) and bool(done[0]) == d_o
        for k, v in obs_o.items():
            tol = 0 if k in ("goal", "fingering") else 3e-4
            np.testing.assert_allclose(obs[k][0], v, atol=tol, err_msg=f"{k} step {step}")
        terms = env.reward_terms
        assert list(terms) == list(ora.reward_terms)
        for name, val in ora.reward_terms.items():
            np.testing.assert_allclose(terms[name][0], val, atol=5e-4, err_msg=f"{name} step {step}")
        np.testing.assert_allclose(rew[0], r_o, atol=2e-3)
        if fingering:
            seen_fingering.add(tuple(obs["fingering"][0]))
    if fingering:
        assert len(seen_fingering) > 1  # the flags follow this hand's notes
    # the device API returns the same compacted row (with the derived position columns)
    env2 = suite.BatchedEnvironment(task, 2, canonical_actions=False, auto_reset=False)
    o_dev = env2.reset_device().cpu().numpy()
    o_np = env2.reset()
    flat = np.concatenate([o_np[k] for k, _ in sorted(env2.layout.items(), key=lambda kv: kv[1][0])], axis=1)
    np.testing.assert_allclose(o_dev.astype(np.float64), flat, atol=1e-7)
    assert o_dev.shape[1] == env2.obs_dim == sum(want.values())
    env.close(); env2.close()
