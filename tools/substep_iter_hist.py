"""Histogram of Newton iterations per substep over envs (config 3 regime, tools/dump_substep_iters.py output) and
what the tail costs a block of 10 envs that meets at every Newton trip.  usage: python tools/substep_iter_hist.py it.npy"""
import sys

import numpy as np

it = np.load(sys.argv[1])  # [substeps, envs]
it = it[20:]  # skip the settling first control steps
print("substeps x envs", it.shape, "mean", it.mean().round(2), "p50 p90 p99 max", np.percentile(it, [50, 90, 99]).tolist(), it.max())
h = np.bincount(it.ravel(), minlength=51)
print("iterations : share of env-substeps")
for k in range(len(h)):
    if h[k]:
        print(f"  {k:3d} : {h[k] / it.size:.4f}")
g = it[:, : it.shape[1] // 10 * 10].reshape(it.shape[0], -1, 10)
print("block of 10: mean of max", g.max(2).mean().round(2), "mean of mean", g.mean(2).mean().round(2),
      "utilisation", (g.mean(2).mean() / g.max(2).mean()).round(3))
for cap in (12, 16, 20, 30):
    c = np.minimum(g, cap)
    print(f"  if no env needed more than {cap}: mean of max {c.max(2).mean():.2f} (time x {c.max(2).mean() / g.max(2).mean():.3f})")
