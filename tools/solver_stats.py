"""Per-step solver statistics of the bench workload (config 2 replay) on the GPU box.

usage: python tools/solver_stats.py [envs] -> per control step: contacts / rows of the last substep, Newton and
line-search iterations, dense-fallback substeps and cross-finger contacts summed over the 10 substeps.
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

import bench  # noqa: E402
from diffusion_piano_b200 import suite  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    task = bench.build_task()
    env = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True)
    table = bench.recorded_actions()
    dev = torch.device("cuda", 0)
    cnt = torch.zeros((B, 8), dtype=torch.int32, device=dev)
    env.reset_device()
    rows = []
    for t in range(len(table)):
        a = torch.from_numpy(np.ascontiguousarray(np.broadcast_to(table[t], (B, 45)))).to(dev)
        env.step_device(a)
        env.engine.get_diag(counters=cnt)
        torch.cuda.synchronize()
        rows.append(cnt[0].cpu().numpy().copy())
    rows = np.asarray(rows)
    print("step ncon nefc newton ls dense lr")
    for t, r in enumerate(rows):
        print(t, r[0], r[1], r[4], r[5], r[6], r[7])
    print("totals over the episode: newton iterations", rows[:, 4].sum(), "line-search", rows[:, 5].sum(),
          "dense-fallback substeps", rows[:, 6].sum(), "of", 10 * len(rows), "; cross-finger contacts per substep mean",
          rows[:, 7].sum() / (10 * len(rows)), "max per step/10", rows[:, 7].max() / 10)
    env.close()


if __name__ == "__main__":
    main()
