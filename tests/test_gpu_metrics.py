"""MidiEvaluationWrapper metrics accumulated by the step on the device, checked against
the reference's own formula (sklearn precision_recall_fscore_support, average="binary", zero_division=1,
per step, then the mean over the episode; reference wrappers/evaluation.py:108-177)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _reference_metrics(truth_rows, pred_rows):
    from sklearn.metrics import precision_recall_fscore_support
    out = []
    for y_true, y_pred in zip(truth_rows, pred_rows):
        p, r, f, _ = precision_recall_fscore_support(y_true=y_true, y_pred=y_pred, average="binary", zero_division=1)
        out.append((p, r, f))
    return np.mean(np.asarray(out), axis=0)


def test_self_actuated_goal_following_policy_scores_one(cuda_lib):
    """reference examples/self_actuated_piano_env.py:84-109: the goal-following policy has reward 0 and
    precision = recall = F1 = 1 for keys and sustain."""
    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.wrappers import MidiEvaluationWrapper
    task = suite.SelfActuatedPiano(music.load("CMajorScaleOneHand"), control_timestep=0.05)
    env = MidiEvaluationWrapper(suite.Environment(task), deque_size=2)
    with pytest.raises(ValueError):
        env.get_musical_metrics()
    hi = env.action_spec().maximum
    for episode in range(2):
        ts = env.reset()
        while not ts.last():
            goal = ts.observation["goal"]
            a = np.zeros(89); a[:88] = np.where(goal[:88] > 0, hi[:88], 0.0); a[88] = goal[88]
            ts = env.step(a)
            assert ts.reward == 0.0
    m = env.get_musical_metrics()
    assert all(v == 1.0 for v in m.values()), m
    env.close()


def test_self_actuated_random_policy_matches_sklearn(cuda_lib):
    """Random key presses (batched, auto-reset): every env's episode metrics equal sklearn's on the recorded
    activations, for two consecutive episodes."""
    import torch
    from diffusion_piano_b200 import music, suite
    task = suite.SelfActuatedPiano(music.load("CMajorScaleTwoHands"), control_timestep=0.05)
    n, T = 6, len(task._notes)
    env = suite.BatchedEnvironment(task, n, canonical_actions=False, auto_reset=True)
    hi = env.action_spec().maximum
    truth = np.zeros((T, 88)); sus_truth = np.zeros((T, 1))
    for t, notes in enumerate(task._notes):
        truth[t, [note.key for note in notes]] = 1.0
        sus_truth[t, 0] = float(task._sustains[t])
    rng = np.random.RandomState(0)
    act = torch.zeros((n, 88), dtype=torch.uint8, device="cuda")
    env.reset()
    for episode in range(2):
        preds = np.zeros((n, T, 88)); sus = np.zeros((n, T, 1))
        for t in range(T):
            a = np.zeros((n, 89), np.float32)
            # press the right keys most of the time, plus a few random wrong ones
            want = truth[t] > 0
            a[:, :88] = np.where((want & (rng.uniform(size=(n, 88)) < 0.8)) | (rng.uniform(size=(n, 88)) < 0.03), hi[:88], 0.0)
            a[:, 88] = (rng.uniform(size=n) < 0.5).astype(np.float32)
            obs, rew, done = env.step(a)
            env.engine.get_diag(activation=act); torch.cuda.synchronize()
            preds[:, t] = act.cpu().numpy(); sus[:, t, 0] = a[:, 88] >= 0.5
            assert done.all() == (t == T - 1)
        m, c = env.episode_metrics()
        assert (c == episode + 1).all()
        for i in range(n):
            np.testing.assert_allclose(m[i, :3], _reference_metrics(truth, preds[i]), atol=2e-6)
            np.testing.assert_allclose(m[i, 3:], _reference_metrics(sus_truth, sus[i]), atol=2e-6)
        env.step(np.zeros((n, 89), np.float32))  # the step after LAST resets every env
    assert 0.0 < env.get_musical_metrics()["f1"] < 1.0
    env.close()


def test_hands_task_metrics_match_sklearn(cuda_lib):
    """Two-hand task: keys pushed by qfrc_applied (as the reference's wrong-press test does) on a schedule,
    metrics against sklearn on the recorded activations."""
    import torch
    from diffusion_piano_b200 import music, suite
    task = suite.PianoWithShadowHands(music.load("CMajorScaleOneHand"), control_timestep=0.25, trim_silence=True)
    T = len(task._notes)
    env = suite.BatchedEnvironment(task, 2, canonical_actions=False, auto_reset=False)
    env.reset()
    truth = np.zeros((T, 88)); sus_truth = np.zeros((T, 1))
    for t, notes in enumerate(task._notes):
        truth[t, [note.key for note in notes]] = 1.0
        sus_truth[t, 0] = float(task._sustains[t])
    act = torch.zeros((2, 88), dtype=torch.uint8, device="cuda")
    f = torch.zeros((2, 140), device="cuda")
    preds = np.zeros((2, T, 88)); sus = np.zeros((2, T, 1))
    lo, hi = env.action_spec().minimum, env.action_spec().maximum
    a = np.stack([0.5 * (lo + hi)] * 2)
    rng = np.random.RandomState(1)
    for t in range(T):
        f.zero_()
        for i in range(2):
            keys = np.flatnonzero((truth[t] > 0) | (rng.uniform(size=88) < 0.05))
            f[i, torch.from_numpy(keys).cuda()] = 3.0
        env.engine.set_qfrc_applied(f); torch.cuda.synchronize()
        a[:, 44] = rng.uniform(size=2)
        obs, rew, done = env.step(a)
        env.engine.get_diag(activation=act); torch.cuda.synchronize()
        preds[:, t] = act.cpu().numpy(); sus[:, t, 0] = a[:, 44] >= 0.5
    assert done.all()
    m, c = env.episode_metrics()
    assert (c == 1).all() and preds.sum() > 0
    for i in range(2):
        np.testing.assert_allclose(m[i, :3], _reference_metrics(truth, preds[i]), atol=2e-6)
        np.testing.assert_allclose(m[i, 3:], _reference_metrics(sus_truth, sus[i]), atol=2e-6)
    env.close()
