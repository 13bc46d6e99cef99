"""Key-press / sustain precision, recall, F1 of the config-2 replay of the reference's recorded policy
(examples/twinkle_twinkle_actions.npy, tutorial.ipynb cells 15, 18-19) on the fp64 CPU oracle - the one
end-to-end physics expectation the reference holds: on the real menagerie hand that action sequence plays the
song. Metrics exactly as MidiEvaluationWrapper (wrappers/evaluation.py:114-177): sklearn binary P/R/F1 per step
with zero_division=1, mean over the episode. Test infrastructure (runs the oracle): tools/ and tests/ only.
usage: python tools/replay_f1.py [--per-step]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def prf(y_true, y_pred):
    tp = float(np.sum((y_true > 0) & (y_pred > 0)))
    fp = float(np.sum((y_true == 0) & (y_pred > 0)))
    fn = float(np.sum((y_true > 0) & (y_pred == 0)))
    p = tp / (tp + fp) if tp + fp > 0 else 1.0
    r = tp / (tp + fn) if tp + fn > 0 else 1.0
    f = 2 * p * r / (p + r) if p + r > 0 else (1.0 if tp + fp + fn == 0 else 0.0)
    return p, r, f


def replay_metrics(per_step=False):
    import bench
    from oracle.task_oracle import PianoWithShadowHandsOracle
    task = bench.build_task()
    acts = bench.recorded_actions()[:158]
    ora = PianoWithShadowHandsOracle(task)
    ora.reset()
    goal = np.asarray(task.song.goal)  # [T, 89] u8
    key, sus, ret = [], [], 0.0
    rows = []
    for t in range(len(acts)):
        _, r, _ = ora.step(ora.canonical_to_spec(acts[t]))
        ret += r
        act = np.asarray(ora.activation, np.float64)
        key.append(prf(goal[t, :88].astype(np.float64), act))
        sus.append(prf(goal[t, 88:89].astype(np.float64), np.array([float(ora.sustain_activation)])))
        rows.append((t, int(goal[t, :88].sum()), int(act.sum()), int(((goal[t, :88] > 0) & (act > 0)).sum())))
    k, s = np.mean(key, 0), np.mean(sus, 0)
    out = dict(precision=k[0], recall=k[1], f1=k[2], sustain_precision=s[0], sustain_recall=s[1], sustain_f1=s[2],
               episode_return=ret, steps=len(acts))
    if per_step:
        for r in rows:
            print("step %3d  goal keys %d  active keys %d  correct %d" % r)
    return out


if __name__ == "__main__":
    m = replay_metrics("--per-step" in sys.argv)
    print({k: round(float(v), 4) for k, v in m.items()})
