"""Counterpart of the reference's `python parallelized_base_v2.py --num_envs N --num_episodes E`
(parallelized_base_v2.py:68-188): PPO v2 on the vectorised two-hands environment, everything on the GPU.
Under torchrun every rank owns `--num_envs` envs and the gradients are all-reduced over NCCL.

    python examples/train_ppo.py --num_envs 4096 --num_episodes 3 --song CrossingFieldCut10s
"""
import argparse
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--num_envs", type=int, default=2)
    ap.add_argument("--num_episodes", type=int, default=20000)
    ap.add_argument("--song", default="CrossingFieldCut10s")
    ap.add_argument("--model_dir", default=None)
    ap.add_argument("--seed", type=int, default=0)
    args = ap.parse_args()
    import torch

    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.ppo import TrainConfig, train

    world, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    pg = None
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
        pg = torch.distributed.group.WORLD
    # task kwargs of parallelized_base_v2.py:28-38, CanonicalSpecWrapper (:44) = canonical_actions
    task = suite.PianoWithShadowHands(music.load(args.song), n_steps_lookahead=1, trim_silence=True,
                                      wrong_press_termination=False, initial_buffer_time=0.0)
    env = suite.BatchedEnvironment(task, args.num_envs, device=local, canonical_actions=True, auto_reset=True)
    print(f"State dimension: {env.obs_dim}")
    print(f"Action dimension: {env.action_dim}")
    cfg = TrainConfig(num_envs=args.num_envs, num_episodes=args.num_episodes, model_dir=args.model_dir, seed=args.seed)
    agent, history = train(env, cfg, process_group=pg)
    if agent.rank == 0:
        print("stats", {k: round(float(v), 5) for k, v in agent.stats().items()})
        print("musical metrics of the last episodes", env.get_musical_metrics())
    agent.close()
    env.close()
    if world > 1:
        torch.distributed.barrier()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
