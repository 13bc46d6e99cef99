"""A few control steps of the compacted schedule on random actions (for ncu launch lists). usage: <envs> <steps> [schedule]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch
from diffusion_piano_b200 import music, suite
n, steps = int(sys.argv[1]), int(sys.argv[2])
sched = sys.argv[3] if len(sys.argv) > 3 else "compacted"
task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)
env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=sched)
g = torch.Generator(device="cuda"); g.manual_seed(1)
env.reset_device()
ms = []
for t in range(steps):
    a = torch.rand((n, 45), device="cuda", generator=g) * 2 - 1
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); env.step_device(a); e1.record()
    torch.cuda.synchronize()
    ms.append(e0.elapsed_time(e1))
print(sched, n, "ms per step:", " ".join(f"{m:.2f}" for m in ms))
