"""Config 1 (SelfActuatedPiano) on the GPU through the C-ABI, against the fp64 oracle and the
reference's own assertions (examples/self_actuated_piano_env.py:82-109;
self_actuated_piano_test.py:78-166)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _task(**kw):
    from diffusion_piano_b200 import music, suite
    args = dict(midi=music.load("TwinkleTwinkleRousseau"), control_timestep=0.01, trim_silence=True)
    args.update(kw)
    return suite.SelfActuatedPiano(**args)


def test_goal_following_policy_scores_zero_and_matches_oracle(cuda_lib):
    from diffusion_piano_b200 import suite
    from oracle.task_oracle import SelfActuatedPianoOracle
    task = _task()
    env = suite.Environment(task)
    ora = SelfActuatedPianoOracle(task)
    spec = env.action_spec()
    assert spec.shape == (89,)
    ts, _ = env.reset(), ora.reset()
    assert ts.first() and ts.reward is None and ts.discount is None
    steps, tp, fp, fn = 0, 0, 0, 0
    while not ts.last():
        goal = ts.observation["goal"]
        a = spec.minimum.copy()
        a[:88] = np.where(goal[:88] > 0, spec.maximum[:88], spec.minimum[:88])
        a[88] = goal[88]
        ts = env.step(a)
        obs_o, r_o, done_o = ora.step(a)
        steps += 1
        assert ts.reward == 0.0 and r_o == 0.0
        assert ts.last() == done_o
        for k, v in obs_o.items():
            np.testing.assert_array_equal(ts.observation[k], v)
    assert steps == len(task._notes) == 786
    env.close()


def test_state_parity_random_ctrl_and_reward_formula(cuda_lib):
    import torch
    from diffusion_piano_b200 import suite
    from oracle.task_oracle import SelfActuatedPianoOracle
    task = _task(n_steps_lookahead=3, control_timestep=0.05)
    n = 8
    benv = suite.BatchedEnvironment(task, n)
    oras = [SelfActuatedPianoOracle(task) for _ in range(n)]
    rng = np.random.RandomState(0)
    hi = benv.engine.action_maximum
    benv.reset()
    qpos = torch.zeros((n, 88), device="cuda")
    qvel = torch.zeros((n, 88), device="cuda")
    for step in range(12):
        a = (rng.uniform(0, 1, (n, 89)) * hi).astype(np.float32)
        obs, rew, done = benv.step(a)
        benv.engine.get_state(qpos=qpos, qvel=qvel)
        q, v = qpos.cpu().numpy(), qvel.cpu().numpy()
        for i, o in enumerate(oras):
            obs_o, r_o, d_o = o.step(a[i].astype(np.float64))
            np.testing.assert_allclose(q[i], o.physics.qpos, atol=1e-5)
            np.testing.assert_allclose(v[i], o.physics.qvel, atol=2e-3, rtol=1e-4)
            assert np.float32(r_o) == np.float32(rew[i])  # bit-exact once activations agree
            for k, val in obs_o.items():
                np.testing.assert_array_equal(obs[k][i], val)
    benv.close()


def test_termination_autoreset_and_xent(cuda_lib):
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import SelfActuatedPianoOracle
    task = suite.SelfActuatedPiano(music.load("CMajorScaleOneHand"), control_timestep=0.5,
                                   reward_type=suite.RewardType.NEGATIVE_XENT)
    T = len(task._notes)
    benv = suite.BatchedEnvironment(task, 3)
    ora = SelfActuatedPianoOracle(task)
    benv.reset()
    a = np.zeros((3, 89), np.float32)
    a[:, 40] = 1.0
    for t in range(T):
        obs, rew, done = benv.step(a)
        _, r_o, d_o = ora.step(a[0].astype(np.float64))
        np.testing.assert_allclose(rew, r_o, rtol=1e-5)
        assert done.all() == d_o == (t == T - 1)
    obs, rew, done = benv.step(a)  # the step after LAST resets (composer semantics)
    assert not done.any() and (rew == 0).all()
    np.testing.assert_array_equal(obs["goal"][0], ora.reset()["goal"])
    benv.close()
