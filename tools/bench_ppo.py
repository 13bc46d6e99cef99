"""Config 4 timing (BASELINE.json configs[3]): 'Crossing Field Cut 10s', the reference's loop of
select_actions -> env step -> PPO v2 update on the batch of that control step, all on the device.
Prints one JSON line per phase breakdown; `python tools/bench_ppo.py --envs 4096 --steps 10`.
Under torchrun every rank owns `--envs` envs and the gradients are all-reduced (NCCL)."""
import argparse
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def measure(envs, steps, warmup, local, world, rank, pg, epochs=10, batch_size=128, impl="product", eager=False,
            graph_scope="update", schedule="auto", gemm="custom"):
    """select_actions -> env step -> PPO update, `steps` control steps after `warmup`; returns (JSON-able dict, agent)."""
    import torch

    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.ppo import PPOAgent

    task = suite.PianoWithShadowHands(music.load("CrossingFieldCut10s"), n_steps_lookahead=1, trim_silence=True)
    env = suite.BatchedEnvironment(task, envs, device=local, canonical_actions=True, auto_reset=True, schedule=schedule)
    if impl == "product":
        agent = PPOAgent(env.obs_dim, env.action_dim, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch_size,
                         ppo_epochs=epochs, device=f"cuda:{local}", seed=0, process_group=pg,
                         use_graphs=not eager, graph_scope=graph_scope, gemm=gemm)
    else:
        sys.path.insert(0, str(ROOT / "tools"))
        from torch_ppo_baseline import TorchPPO

        torch.manual_seed(0)
        agent = TorchPPO(env.obs_dim, env.action_dim, batch_size=batch_size, ppo_epochs=epochs,
                         device=f"cuda:{local}", python_gae=impl == "torch")
        agent.kernel_launches = 0
        agent.stats = lambda: {}
    n = envs
    states = torch.empty((n, env.obs_dim), dtype=torch.float32, device=env.device)
    obs = env.reset_device()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
    launches0 = 0
    for it in range(warmup + steps):
        if it == warmup:
            if pg is not None:
                torch.distributed.barrier()
            torch.cuda.synchronize()
            launches0 = agent.kernel_launches + env.engine.launch_count
        e = ev[it - warmup] if it >= warmup else None
        states.copy_(obs)
        if e: e[0].record()
        actions, logp = agent.select_actions(states)
        if e: e[1].record()
        obs, rewards, dones, _, _ = env.step_device(actions)
        if e: e[2].record()
        agent.update(states, actions, rewards, logp, obs, dones)
        if e: e[3].record()
    torch.cuda.synchronize()
    launches = agent.kernel_launches + env.engine.launch_count - launches0
    act = sum(a[0].elapsed_time(a[1]) for a in ev) / steps
    step = sum(a[1].elapsed_time(a[2]) for a in ev) / steps
    upd = sum(a[2].elapsed_time(a[3]) for a in ev) / steps
    total = torch.tensor([ev[0][0].elapsed_time(ev[-1][3])], device=env.device)
    if pg is not None:
        torch.distributed.all_reduce(total, op=torch.distributed.ReduceOp.MAX)
    total_ms = float(total.item())
    mbs = epochs * ((n + batch_size - 1) // batch_size)
    line = {
        "metric": "PPO v2 rollout+update env steps/sec (config 4)", "value": n * world * steps / (total_ms * 1e-3),
        "unit": "env-steps/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": total_ms / steps, "phases_ms": {"select_actions": act, "env_step": step, "update": upd},
        "minibatches_per_update": mbs, "us_per_minibatch": 1e3 * upd / mbs, "gpu_launches": launches,
        "impl": impl, "graphs": not eager, "schedule": env.engine.get_schedule(), "stats": agent.stats(),
        "config": {"workload": "PianoWithShadowHands 'Crossing Field Cut 10s', ppo_v2 update every control step",
                   "envs_per_gpu": n, "batch_size": batch_size, "ppo_epochs": epochs,
                   "obs_dim": env.obs_dim, "parallelism": f"env-sharded x{world}, gradient all-reduce per minibatch"}}
    return line, agent


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--envs", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--epochs", type=int, default=10)
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--eager", action="store_true", help="no CUDA graphs (launch-bound comparison)")
    ap.add_argument("--impl", default="product", choices=["product", "torch", "torch-hostgae"],
                    help="torch: plain-PyTorch restatement of ppo_v2.update (tools/torch_ppo_baseline.py) on the same env")
    ap.add_argument("--graph-scope", default="update", choices=["update", "minibatch"])
    ap.add_argument("--schedule", default="auto", choices=["auto", "lockstep", "divergent", "compacted"])
    ap.add_argument("--gemm", default="custom", choices=["custom", "cublas"])
    ap.add_argument("--blas", default=None, help="torch.backends.cuda.preferred_blas_library (cublas / cublaslt)")
    args = ap.parse_args()
    import faulthandler

    import torch

    if os.environ.get("PPO_BENCH_DUMP_AFTER"):  # where is it stuck: Python stacks of every thread after N seconds
        faulthandler.dump_traceback_later(int(os.environ["PPO_BENCH_DUMP_AFTER"]), exit=True)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if args.blas:
        torch.backends.cuda.preferred_blas_library(args.blas)
    pg = None
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
        pg = torch.distributed.group.WORLD
    line, agent = measure(args.envs, args.steps, args.warmup, local, world, rank, pg, epochs=args.epochs,
                          batch_size=args.batch_size, impl=args.impl, eager=args.eager, graph_scope=args.graph_scope,
                          schedule=args.schedule, gemm=args.gemm)
    if rank == 0:
        print(json.dumps(line))
    if pg is not None:
        # destroy_process_group() blocks when CUDA graphs that captured NCCL kernels were replayed on its
        # communicators (seen on 2 x B200): release the graphs, meet at a barrier and leave without it
        if hasattr(agent, "close"):
            agent.close()
        torch.distributed.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
