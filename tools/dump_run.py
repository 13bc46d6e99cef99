"""Random-action run under a launch shape, results to an .npz (to compare two builds of libpiano_step.so bit for bit:
PIANO_STEP_LIB=build/a.so python tools/dump_run.py 333 8 a.npz lockstep; ... b.npz; python tools/dump_run.py --cmp a.npz b.npz)"""
import sys
from pathlib import Path

import numpy as np

if sys.argv[1] == "--cmp":
    a, b = np.load(sys.argv[2]), np.load(sys.argv[3])
    worst = {k: float(np.abs(a[k].astype(np.float64) - b[k]).max()) for k in a.files}
    same = all(np.array_equal(a[k], b[k]) for k in a.files)
    print("identical" if same else f"DIFFERENT {worst}")
    sys.exit(0)
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from diffusion_piano_b200 import music, suite  # noqa: E402

n, steps, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
sched = sys.argv[4] if len(sys.argv) > 4 else "lockstep"
task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)
env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=sched)
g = torch.Generator(device="cuda"); g.manual_seed(1)
env.reset_device()
qpos, cnt = torch.zeros((n, 140), device="cuda"), torch.zeros((n, 8), dtype=torch.int32, device="cuda")
rows = {"obs": [], "rew": [], "qpos": [], "cnt": []}
for t in range(steps):
    a = torch.rand((n, 45), device="cuda", generator=g) * 2 - 1
    obs, rew, done, terms, _ = env.step_device(a)
    env.engine.get_state(qpos=qpos); env.engine.get_diag(counters=cnt)
    torch.cuda.synchronize()
    for k, v in (("obs", obs), ("rew", rew), ("qpos", qpos), ("cnt", cnt)):
        rows[k].append(v.cpu().numpy().copy())
np.savez(out, **{k: np.stack(v) for k, v in rows.items()})
