"""Config 3 regime (uniform random canonical actions): ms per control step and solver statistics averaged over
the envs, per block of steps. usage: python tools/random_action_stats.py [envs] [steps] [sigma] [schedule] [group]
sigma > 0: actions ~ clip(N(0, sigma)) instead of U(-1, 1) (what an untrained ppo_v2 policy emits, sigma = e^-1)."""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from diffusion_piano_b200 import music, suite  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 60
    sigma = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
    schedule = sys.argv[4] if len(sys.argv) > 4 else "auto"
    group = int(sys.argv[5]) if len(sys.argv) > 5 else 1  # every action row is shared by `group` consecutive envs
    task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)
    env = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True, schedule=schedule)
    dev = torch.device("cuda", 0)
    cnt = torch.zeros((B, 8), dtype=torch.int32, device=dev)
    g = torch.Generator(device=dev)
    g.manual_seed(0)
    env.reset_device()
    print("step  ms   ncon nefc newton ls dense lowrank   (means over envs; max newton, max ncon)")
    for t in range(steps):
        if sigma > 0:
            a = (sigma * torch.randn((B, 45), device=dev, generator=g)).clamp_(-1, 1)
        else:
            a = torch.rand((B // group, 45), device=dev, generator=g) * 2 - 1
            if group > 1:
                a = a.repeat_interleave(group, 0).contiguous()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        env.step_device(a)
        e1.record()
        env.engine.get_diag(counters=cnt)
        torch.cuda.synchronize()
        c = cnt.float()
        m = c.mean(0).cpu().numpy()
        if t % 5 == 0 or t == steps - 1:
            print(f"{t:4d} {e0.elapsed_time(e1):6.2f} {m[0]:5.2f} {m[1]:6.1f} {m[4]:6.1f} {m[5]:6.1f} {m[6]:5.2f} {m[7]:6.1f}"
                  f"   {int(c[:, 4].max())} {int(c[:, 0].max())}")
    print("schedule", env.engine.get_schedule())
    env.close()


if __name__ == "__main__":
    main()
