"""What makes an env's Newton trips slow? Diagnostic build (tools/build_variant.sh sp_timing -DPP_SP_TIMING) of the
compacted schedule: per env and SUBSTEP (control_timestep = physics timestep) the cycles its trips took without the
block meetings, beside its contact / coupling / iteration counts.  usage: PIANO_STEP_LIB=build/sp_timing.so python
tools/trip_cost_model.py <envs> <control steps> <out.npy> -> int32 [substeps, envs, 8] (the counters array)"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import torch  # noqa: E402

from diffusion_piano_b200 import music, suite  # noqa: E402

B, steps, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True,
                                  control_timestep=0.005)
env = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True, schedule="compacted")
g = torch.Generator(device="cuda")
g.manual_seed(0)
env.reset_device()
cnt = torch.zeros((B, 8), dtype=torch.int32, device="cuda")
rows = []
for t in range(steps):
    a = torch.rand((B, 45), device="cuda", generator=g) * 2 - 1
    for _ in range(10):
        env.step_device(a)
        env.engine.get_diag(counters=cnt)
        rows.append(cnt.cpu().numpy().copy())
np.save(out, np.asarray(rows))
