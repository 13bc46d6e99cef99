"""Where the time of a host-buffer step goes (pp_step_host, the `e2e` figure of bench.py): CUDA-event timeline
of config 2 at B = 4096 with pinned host buffers, mean over the timed calls.
usage: python tools/step_host_timeline.py [envs] [calls]   (PP_HOST_CHUNKS=k to change the number of pieces)"""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402


def main():
    import torch
    from diffusion_piano_b200 import suite
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    calls = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    task = bench.build_task()
    env = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True, schedule="lockstep")
    table = bench.recorded_actions()
    pin = lambda shape, dt: torch.zeros(shape, dtype=dt).pin_memory().numpy()
    h_act = [np.ascontiguousarray(np.broadcast_to(table[i % len(table)], (B, 45))) for i in range(calls + 5)]
    h_act = [torch.from_numpy(a.copy()).pin_memory().numpy() for a in h_act]
    h_obs, h_rew, h_done = pin((B, env.obs_dim), torch.float32), pin((B,), torch.float32), pin((B,), torch.uint8)
    h_terms, h_disc = pin((B, 5), torch.float32), pin((B,), torch.float32)
    for i in range(5):
        env.engine.step_host(h_act[i], h_obs, h_rew, h_done, h_terms, h_disc)
    env.engine.host_timeline(True)
    rows, wall = [], []
    for i in range(calls):
        t0 = time.perf_counter()
        env.engine.step_host(h_act[5 + i], h_obs, h_rew, h_done, h_terms, h_disc)
        wall.append(1e3 * (time.perf_counter() - t0))
        h2d, pieces = env.engine.host_timeline(True)
        rows.append([h2d] + [x for p in pieces for x in p])
    m = np.mean(rows, 0)
    print(f"B = {B}, {calls} calls, obs_dim {env.obs_dim}: host wall per call {np.mean(wall):.3f} ms "
          f"(= {B / np.mean(wall) * 1e3:.0f} env-steps/s); D2H bytes per call {B * (env.obs_dim + 7) * 4 + B}")
    print(f"  actions on the device at {m[0]:.3f} ms")
    for k in range((len(m) - 1) // 2):
        print(f"  piece {k}: step done at {m[1 + 2 * k]:.3f} ms, results on the host at {m[2 + 2 * k]:.3f} ms "
              f"(copy {m[2 + 2 * k] - m[1 + 2 * k]:.3f} ms after its step)")
    print(f"  last event at {m[1:].max():.3f} ms; host wall - last event = {np.mean(wall) - m[1:].max():.3f} ms "
          f"(launch + synchronise + return)")
    env.close()


if __name__ == "__main__":
    main()
