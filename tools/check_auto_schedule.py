"""The handle's launch-shape auto-tuner on the config-2 replay (every env identical): it must keep 2 envs per block.
usage (GPU box): python tools/check_auto_schedule.py"""
import sys; from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import numpy as np, torch, bench
from diffusion_piano_b200 import suite
task = bench.build_task(); table = bench.recorded_actions()
env = suite.BatchedEnvironment(task, 4096, canonical_actions=True, auto_reset=True)
env.reset_device()
for t in range(24):
    a = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(table[t], (4096, 45)))).cuda()
    env.step_device(a)
torch.cuda.synchronize(); print("config 2 replay, auto:", env.engine.get_schedule())
