"""Config-2 replay (every env identical) under each launch shape: ms per control step over steps 5..54.
usage (GPU box): python tools/replay_schedules.py [envs]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch, bench
from diffusion_piano_b200 import suite
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
task = bench.build_task(); table = bench.recorded_actions()
for sched in ("lockstep", "divergent", "compacted"):
    env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=sched)
    env.reset_device()
    ms = []
    for t in range(55):
        a = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(table[t], (n, 45)))).cuda()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); env.step_device(a); e1.record(); torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    m = float(np.mean(ms[5:]))
    print(f"{sched:10s} {n} envs: {m:.3f} ms per step = {n / m:.0f} k env-steps/s")
    env.close()
