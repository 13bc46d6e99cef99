"""`legacy_step=False` (reference suite.load(..., legacy_step=) -> composer.Environment -> dm_control Physics.step):
the physics is stepped by plain mj_step, with no trailing mj_step1, so everything the task derives from positions
and velocities after the step - fingertip sites, key geoms, contacts (forearm reward), actuator velocities (energy
reward), a hand's root-body position - is what the LAST mj_step's forward pass computed, i.e. one substep old.
qpos / qvel and everything read from them directly (key activation, joint observations) are the same in both modes.
The CUDA step against the fp64 oracle run in the same mode, and against the CUDA step in the default mode."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(task, legacy, n=2):
    import torch
    from diffusion_piano_b200 import suite
    from oracle.task_oracle import PianoWithShadowHandsOracle
    env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=False, legacy_step=legacy)
    ora = PianoWithShadowHandsOracle(task, legacy_step=legacy)
    env.reset(); ora.reset()
    return env, ora


def _sync(env, ora, bufs):
    """env 0 of the batch takes the oracle's state (qpos, qvel, warm start) so that one control step is compared."""
    import torch
    qpos, qvel, warm = bufs
    p = ora.physics
    env.engine.get_state(qpos=qpos, qvel=qvel, warm=warm)
    for dst, src in ((qpos, p.qpos), (qvel, p.qvel), (warm, p.qacc_warmstart)):
        dst[:] = torch.from_numpy(src.astype(np.float32)).cuda()
    env.engine.set_state(qpos=qpos, qvel=qvel, warm=warm)
    torch.cuda.synchronize()


@pytest.mark.parametrize("fingering", [True, False], ids=["fingering", "optimal-transport"])
def test_non_legacy_step_reads_the_last_substeps_forward_pass(fingering):
    import torch
    import bench
    from diffusion_piano_b200 import music, suite
    midi = music.load("TwinkleTwinkleRousseau")
    task = suite.PianoWithShadowHands(midi, n_steps_lookahead=1, trim_silence=True,
                                      disable_fingering_reward=not fingering)
    acts = bench.recorded_actions()
    env_n, ora_n = _pair(task, legacy=False)
    env_l, ora_l = _pair(task, legacy=True)
    z = lambda *s: torch.zeros(s, device="cuda")
    bufs = (z(2, 140), z(2, 140), z(2, 140))
    st_n, st_l = (z(2, 140), z(2, 140)), (z(2, 140), z(2, 140))
    sx_n, sx_l, av_n = z(2, 10, 3), z(2, 10, 3), z(2, 44)
    name = "fingering_reward" if fingering else "ot_fingering_reward"
    site_gap, fing_gap, energy_gap = [], [], []
    rng = np.random.RandomState(3)
    for t in range(40):
        # the recorded policy with a random kick every few steps: fingers near their keys AND moving fast
        a = acts[t].copy()
        if t % 4 == 3:
            a[:44] = np.clip(a[:44] + rng.uniform(-0.6, 0.6, 44), -1, 1)
        _sync(env_n, ora_n, bufs)
        _sync(env_l, ora_n, bufs)  # both modes start the step from the same state
        A = np.stack([a, a]).astype(np.float32)
        obs_n, rew_n, done_n = env_n.step(A)
        obs_l, rew_l, done_l = env_l.step(A)
        obs_o, r_o, d_o = ora_n.step(ora_n.canonical_to_spec(a))
        env_n.engine.get_state(qpos=st_n[0], qvel=st_n[1]); env_l.engine.get_state(qpos=st_l[0], qvel=st_l[1])
        env_n.engine.get_diag(site_xpos=sx_n, actuator_velocity=av_n); env_l.engine.get_diag(site_xpos=sx_l)
        torch.cuda.synchronize()
        # 1. the physics does not know about the flag: same qpos / qvel, bit for bit
        assert torch.equal(st_n[0], st_l[0]) and torch.equal(st_n[1], st_l[1])
        for k in ("piano/state", "rh_shadow_hand/joints_pos", "lh_shadow_hand/joints_pos", "goal"):
            np.testing.assert_array_equal(obs_n[k], obs_l[k])
        # 2. against the oracle in the same mode: derived quantities one substep old
        p = ora_n.physics
        np.testing.assert_allclose(sx_n[0].cpu().numpy(), p.site_xpos, atol=1e-4, err_msg=f"site_xpos, step {t}")
        tn = {k: float(v[0]) for k, v in env_n.reward_terms.items()}
        for k, v in ora_n.reward_terms.items():
            tol = 2e-3 if k == "energy_reward" else 1e-3
            assert abs(tn[k] - v) <= tol + 1e-3 * abs(v), f"{k} step {t}: gpu {tn[k]} oracle {v}"
        assert abs(rew_n[0] - r_o) <= 5e-3
        assert bool(done_n[0]) == d_o
        # 3. and they do differ from the default mode's
        site_gap.append(float((sx_n[0] - sx_l[0]).abs().max()))
        tl = {k: float(v[0]) for k, v in env_l.reward_terms.items()}
        fing_gap.append(abs(tn[name] - tl[name])); energy_gap.append(abs(tn["energy_reward"] - tl["energy_reward"]))
    print(f"legacy_step False against True over 40 steps: fingertip sites differ by up to {max(site_gap):.2e} m, "
          f"{name} by up to {max(fing_gap):.2e}, energy_reward by up to {max(energy_gap):.2e}")
    assert max(site_gap) > 5e-4 and max(energy_gap) > 1e-4 and max(fing_gap) > 1e-5
    env_n.close(); env_l.close()


def test_non_legacy_reset_is_the_same_reset():
    """A reset runs a full forward pass at the reset state in both modes (composer calls physics.forward())."""
    import torch
    from diffusion_piano_b200 import music, suite
    task = suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), n_steps_lookahead=2)
    outs = []
    for legacy in (True, False):
        env = suite.BatchedEnvironment(task, 2, legacy_step=legacy)
        obs = env.reset()
        sx = torch.zeros((2, 10, 3), device="cuda")
        env.engine.get_diag(site_xpos=sx)
        torch.cuda.synchronize()
        outs.append((obs, sx.cpu().numpy()))
        env.close()
    for k in outs[0][0]:
        np.testing.assert_array_equal(outs[0][0][k], outs[1][0][k])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])


def test_one_hand_position_observable_is_one_substep_old():
    """`<hand>/position` = xpos of the hand's root body (hands/base.py:110-113): under legacy_step=False it lags the
    forearm joints it is built from by one substep."""
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import PianoWithShadowHandsOracle
    task = suite.PianoWithOneShadowHand(music.load("CMajorScaleTwoHands"), hand_side=suite.HandSide.RIGHT,
                                        n_steps_lookahead=1)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=True, auto_reset=False, legacy_step=False)
    ora = PianoWithShadowHandsOracle(task, legacy_step=False)
    env.reset(); ora.reset()
    red = task.reduced_spec()
    rng = np.random.RandomState(5)
    lag = []
    for t in range(6):
        a = rng.uniform(-1, 1, len(red[0]))
        full = np.zeros(45)
        full[np.asarray(red[0])] = a
        obs, rew, done = env.step(np.stack([a, a]))
        obs_o, r_o, _ = ora.step(ora.canonical_to_spec(full.astype(np.float32)))
        np.testing.assert_allclose(obs["rh_shadow_hand/position"][0], obs_o["rh_shadow_hand/position"], atol=2e-4)
        p = ora.physics  # the position an up-to-date forward pass would give
        p.forward_position(False)
        body = task.model.names["body"].index("rh_shadow_hand/rh_forearm")
        lag.append(float(np.abs(p.xpos[body] - obs_o["rh_shadow_hand/position"]).max()))
    print(f"one-hand position observable lags the joints by up to {max(lag):.2e} m")
    assert max(lag) > 1e-4
    env.close()
