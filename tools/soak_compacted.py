"""Soak run: many control steps of random actions (auto-resetting episodes) under a launch shape; every observation
and reward must stay finite and the musical metrics sane.  usage: python tools/soak_compacted.py <envs> <steps> [schedule]"""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from diffusion_piano_b200 import music, suite  # noqa: E402

n, steps = int(sys.argv[1]), int(sys.argv[2])
sched = sys.argv[3] if len(sys.argv) > 3 else "compacted"
task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True,
                                  wrong_press_termination=True)
env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=sched)
g = torch.Generator(device="cuda"); g.manual_seed(3)
env.reset_device()
bad = torch.zeros((), dtype=torch.int64, device="cuda")
dones = torch.zeros((), dtype=torch.int64, device="cuda")
rsum = torch.zeros((), dtype=torch.float64, device="cuda")
cnt = torch.zeros((n, 8), dtype=torch.int32, device="cuda")
t0 = time.time()
for t in range(steps):
    a = torch.rand((n, 45), device="cuda", generator=g) * 2 - 1
    if t % 7 == 0:
        a[:, :44] = a[:, :44].sign()  # slam every actuator to a bound now and then
    obs, rew, done, terms, disc = env.step_device(a)
    bad += (~torch.isfinite(obs)).sum() + (~torch.isfinite(rew)).sum()
    dones += done.sum()
    rsum += rew.double().sum()
env.engine.get_diag(counters=cnt)
torch.cuda.synchronize()
dt = time.time() - t0
print(f"{sched} {n} envs x {steps} steps in {dt:.1f} s ({n * steps / dt / 1e3:.0f} k env-steps/s wall): non-finite values {int(bad)}, "
      f"episodes ended {int(dones)}, mean reward {float(rsum) / (n * steps):.4f}, dropped contacts in the last step {int(cnt[:, 3].sum())}, "
      f"max Newton iterations of a substep {int(cnt[:, 2].max())}")
assert int(bad) == 0
