"""Task options of `PianoTask` that change the scene (reference suite/tasks/base.py:96-107,149-197):
`attachment_yaw` (forearms rotated about world z, opposite signs for the two hands) and `forearm_dofs` (which
forearm joints the hands have, shadow_hand.py:41-69,272-311; subsets of the default pair - the missing slide is
locked and leaves the action / observation specs). The CUDA step against the fp64 oracle on the same tables."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kw", [dict(attachment_yaw=6.0), dict(attachment_yaw=-10.0, gravity_compensation=True),
                                dict(forearm_dofs=("forearm_tx",)), dict(forearm_dofs=("forearm_ty",)),
                                dict(forearm_dofs=())], ids=lambda kw: "-".join(f"{k}={v}" for k, v in kw.items()))
def test_scene_options_parity(kw):
    import torch
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import PianoWithShadowHandsOracle

    midi = music.load("TwinkleTwinkleRousseau")
    task = suite.PianoWithShadowHands(midi, n_steps_lookahead=1, trim_silence=True, **kw)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=True, auto_reset=False)
    n_fore = len(kw.get("forearm_dofs", ("forearm_tx", "forearm_ty")))
    # reference shadow_hand_test.py:81-99: len(joints) = NQ + n_forearm_dofs, len(actuators) = NU + n_forearm_dofs
    assert env.action_spec().shape == (2 * (20 + n_fore) + 1,)
    spec = env.observation_spec()
    assert spec["rh_shadow_hand/joints_pos"].shape == (24 + n_fore,) == spec["lh_shadow_hand/joints_pos"].shape
    ora = PianoWithShadowHandsOracle(task)
    red = task.reduced_spec()
    act_keep = red[0] if red else list(range(45))
    obs = env.reset()
    obs_o = ora.reset()
    sx = torch.zeros((2, 10, 3), device="cuda")
    env.engine.get_diag(site_xpos=sx)
    torch.cuda.synchronize()
    np.testing.assert_allclose(sx[0].cpu().numpy(), ora.physics.site_xpos, atol=2e-6)  # the rotated / fixed pose
    if "attachment_yaw" in kw:  # fingertips swing towards the other hand for a positive yaw, mirror-symmetrically
        plain = PianoWithShadowHandsOracle(suite.PianoWithShadowHands(midi, trim_silence=True))  # keeps its buffers alive
        dy = ora.physics.site_xpos[:, 1] - plain.physics.site_xpos[:, 1]
        assert (np.sign(dy[:5]) == -np.sign(kw["attachment_yaw"])).all() and np.allclose(dy[:5], -dy[5:], atol=1e-9)
    rng = np.random.RandomState(0)
    qpos = torch.zeros((2, 140), device="cuda")
    cnt = torch.zeros((2, 8), dtype=torch.int32, device="cuda")
    for step in range(6):
        a = rng.uniform(-1, 1, len(act_keep))
        full = np.zeros(45)
        full[act_keep] = a
        obs, rew, done = env.step(np.stack([a, a]))
        obs_o, r_o, _ = ora.step(ora.canonical_to_spec(full.astype(np.float32)))
        env.engine.get_diag(counters=cnt)
        torch.cuda.synchronize()
        assert int(cnt[0, 3]) == 0, "the step dropped contacts (kernel caps): the comparison would be meaningless"
        for h in ("rh", "lh"):
            k = f"{h}_shadow_hand/joints_pos"
            lanes = [d - 88 - (0 if h == "rh" else 26) for d in task.model.locked_dofs]
            drop = {(l - 2 if l >= 2 else 24 + l) for l in lanes if 0 <= l < 26}
            keep = [j for j in range(26) if j not in drop]
            np.testing.assert_allclose(obs[k][0], obs_o[k][keep], atol=3e-4, err_msg=f"{k} step {step}")
        np.testing.assert_allclose(obs["piano/state"][0], obs_o["piano/state"], atol=3e-4)
        np.testing.assert_allclose(rew[0], r_o, atol=5e-3)
        if task.model.locked_dofs:
            env.engine.get_state(qpos=qpos)
            torch.cuda.synchronize()
            assert qpos[:, task.model.locked_dofs].abs().max().item() < 1e-6  # a missing slide does not move
    env.close()


def test_unbuilt_forearm_dofs_raise():
    from diffusion_piano_b200 import music, suite
    with pytest.raises(NotImplementedError, match="forearm_roll"):
        suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), forearm_dofs=("forearm_tx", "forearm_roll"))
