"""Compacted schedule against the lockstep one on random actions: first differing step / field, then timings."""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch
from diffusion_piano_b200 import music, suite

n, steps = int(sys.argv[1]), int(sys.argv[2])
task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)


def run(schedule, time_it=False):
    env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=schedule)
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    env.reset_device()
    out, ms = [], []
    qpos, cnt = torch.zeros((n, 140), device="cuda"), torch.zeros((n, 8), dtype=torch.int32, device="cuda")
    for t in range(steps):
        a = torch.rand((n, 45), device="cuda", generator=g) * 2 - 1
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        obs, rew, done, terms, _ = env.step_device(a)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
        env.engine.get_state(qpos=qpos); env.engine.get_diag(counters=cnt)
        torch.cuda.synchronize()
        if not time_it:
            out.append((obs.cpu().numpy().copy(), rew.cpu().numpy().copy(), terms.cpu().numpy().copy(), qpos.cpu().numpy().copy(), cnt.cpu().numpy().copy()))
    env.close()
    return out, ms


ref, ms0 = run("lockstep"); print("lockstep done", flush=True)
new, ms1 = run("compacted"); print("compacted done", flush=True)
bad = False
for t, (r, c) in enumerate(zip(ref, new)):
    for name, x, y in zip(("obs", "reward", "terms", "qpos", "counters"), r, c):
        if not np.array_equal(x, y):
            d = np.argwhere(x != y)
            print(f"step {t} {name}: {len(d)} entries differ, first {d[0]}, max |diff| {np.abs(x.astype(np.float64) - y).max():.3e}; envs {np.unique(d[:, 0])[:8]}")
            bad = True
    if bad:
        break
print("identical" if not bad else "DIFFERENT")
_, ms2 = run("divergent", True)
f = lambda m: float(np.mean(m[5:]))
print(f"n {n}: lockstep {f(ms0):.3f} ms  divergent {f(ms2):.3f} ms  compacted {f(ms1):.3f} ms per control step")
