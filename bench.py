"""bench.py - env-steps/s of the batched PianoWithShadowHands step (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            # product arm (CUDA step, C-ABI)
  python bench.py --impl reference --gpus N --steps K ...  # CPU arm: the fp64 oracle port, all host cores

Workload (BASELINE.json configs[1]): PianoWithShadowHands on
TwinkleTwinkleRousseau, task kwargs of reference tutorial.ipynb cell 15 (trim_silence,
control_timestep 0.05 = 10 substeps, gravity_compensation, n_steps_lookahead 10), every env
replaying reference examples/twinkle_twinkle_actions.npy row t at step t, 4096 envs per GPU.
A "step" is one control step of every env of the batch (one kernel launch per GPU).

One JSON line is printed by rank 0. `value` = env-steps/s over all ranks with actions already
in HBM (device timing, CUDA events on the launching stream, max over ranks); `e2e` = the SAME
control steps of the episode through pp_step_host with HOST buffers (H2D of actions, D2H of
obs/reward/done inside the timed region). MuJoCo itself is absent from this box ("unavailable on
this box"), so the CPU arm is the fp64 restatement in oracle/ (kind "port"), OpenMP over envs.
`extra` carries the other BASELINE configs, each timed with its own CUDA events after the headline
(--extras 3,5,4; "" to skip): config3 (random actions, B = 16384), config5 (contact stress,
B = 65536), config4_ppo (rollout + PPO v2 update with the gradient all-reduce across ranks).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "env steps/sec, PianoWithShadowHands B=4096 per GPU"
UNIT = "env-steps/s"
L2_FLUSH_BYTES = 256 << 20


def bytes_per_env_step(obs_dim: int) -> int:
    """Algorithmic HBM bytes of one env step: fp32 state in+out, action, obs, scalars."""
    read = 3 * 140 * 4 + 45 * 4 + 4
    write = 3 * 140 * 4 + obs_dim * 4 + 7 * 4 + 1 + 88
    return read + write


def build_task():
    from diffusion_piano_b200 import music, suite

    return suite.PianoWithShadowHands(
        midi=music.load("TwinkleTwinkleRousseau"), trim_silence=True, control_timestep=0.05,
        gravity_compensation=True, primitive_fingertip_collisions=False, reduced_action_space=False,
        n_steps_lookahead=10, change_color_on_activation=True)


def recorded_actions() -> np.ndarray:
    d = np.load(ROOT / "diffusion-piano_b200" / "data" / "recorded_actions.npz")
    a = d["twinkle_twinkle"].astype(np.float32)  # (158, 45), canonical [-1, 1]
    # an env whose previous step was LAST spends the next call resetting (composer semantics):
    # row 158 is that filler call, then the replay starts over
    return np.concatenate([a, np.zeros((1, 45), np.float32)], 0)


def workload_config(envs_per_gpu: int, n_gpus: int, task) -> dict:
    return {"workload": "PianoWithShadowHands TwinkleTwinkleRousseau, twinkle_twinkle_actions.npy replay",
            "envs_per_gpu": envs_per_gpu, "global_envs": envs_per_gpu * n_gpus, "substeps_per_step": 10,
            "n_steps_lookahead": 10, "obs_dim": 1130, "episode_frames": len(task._notes),
            "gravity_compensation": True, "schedule": "lockstep (2 envs per block; all envs replay one action table)", "parallelism": f"env-sharded x{n_gpus}, no collective in the step",
            "l2": f"flushed between timed steps ({L2_FLUSH_BYTES >> 20} MiB write)"}


# ------------------------------------------------------------------ CPU arm (oracle port)
class CpuPort:
    """n independent fp64 oracle envs stepped with OpenMP over envs (oracle/piano_oracle.c).
    Physics only: the task epilogue (rewards, observation gather) is < 1 % of a CPU env step.
    sparse=True (the timed baseline): po_step_sparse - sparse constraint rows, one dense block over the coupled
    dofs, no allocation in the step; same results as the dense clarity-first po_step to 1e-12
    (tests/test_scene_and_oracle.py) and ~8x faster."""

    def __init__(self, task, n_envs: int, sparse: bool = True):
        from oracle import oracle_physics as op

        self.op = op
        self.n = n_envs
        self.envs = [op.OraclePhysics(task.model) for _ in range(n_envs)]
        self.ptrs = (ctypes.POINTER(op._Data) * n_envs)(*[e._d for e in self.envs])
        self.mp = self.envs[0]._mp
        a = task.model.a
        self.lo, self.hi = a["act_ctrlrange"][:, 0].copy(), a["act_ctrlrange"][:, 1].copy()
        self.nsub = task.physics_steps_per_control_step
        self.batch = op.lib().po_step_batch_sparse if sparse else op.lib().po_step_batch
        # all host cores, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1)
        op.lib().po_set_num_threads(os.cpu_count() or 1)

    def step(self, canonical_action: np.ndarray) -> None:
        ctrl = self.lo + 0.5 * (canonical_action[:44].astype(np.float64) + 1.0) * (self.hi - self.lo)
        for e in self.envs:
            e.ctrl[:] = ctrl
        self.batch(self.mp, self.ptrs, self.n, self.nsub)


def time_cpu_port(task, actions: np.ndarray, n_envs: int, steps: int, warmup: int, sparse: bool = True) -> float:
    """env-steps/s of the oracle port on the host cores."""
    port = CpuPort(task, n_envs, sparse)
    for i in range(warmup):
        port.step(actions[i % len(actions)])
    t0 = time.perf_counter()
    for i in range(steps):
        port.step(actions[(warmup + i) % len(actions)])
    return n_envs * steps / (time.perf_counter() - t0)


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    task = build_task()
    cores = os.cpu_count() or 1
    n_envs = 8 * cores
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    t0 = time.perf_counter()
    v = time_cpu_port(task, recorded_actions(), n_envs, args.steps, max(1, min(args.warmup, 2)))
    sample = (f"{n_envs} envs x {args.steps} control steps of the workload (10 fp64 substeps each), "
              f"OpenMP over envs, po_step_sparse (block-structured physics restatement, no allocation in the step); "
              f"MuJoCo itself unavailable on this box")
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_envs / v,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args.envs, args.gpus, task),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ clocks during the timed region
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            f = tempfile.NamedTemporaryFile("w", suffix=".csv", delete=False)
            self.path = f.name
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=f,
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for row in Path(self.path).read_text().splitlines():
                p = [x.strip() for x in row.split(",")]
                if len(p) < 9:
                    continue
                sm.append(float(p[1])); mx.append(float(p[2]))
                for name, val in zip(names, p[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


# ------------------------------------------------------------------ product arm
def run_product(args) -> None:
    import torch
    import torch.distributed as dist

    from diffusion_piano_b200 import sharding, suite

    shard = sharding.shard_from_env(args.envs)  # rank r owns global envs [r*B, (r+1)*B): no collective in the step
    rank, world, local = shard.rank, shard.world_size, shard.local_rank
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product arm has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    task = build_task()
    B, K, Wm = args.envs, args.steps, args.warmup
    # every env replays the same recorded action row: the "lockstep" launch shape (2 envs per block) - what the
    # handle's auto-tuner (pp_set_schedule) also picks for this workload, fixed here so that its 12 timing steps do
    # not fall inside the timed region
    env = suite.BatchedEnvironment(task, B, device=local, canonical_actions=True, auto_reset=True,
                                   schedule="lockstep")
    table = recorded_actions()
    period = len(table)
    # inputs resident in HBM before the timed region: one (B,45) tensor per call
    acts = [torch.from_numpy(np.ascontiguousarray(np.broadcast_to(table[i % period], (B, 45)))).to(dev)
            for i in range(Wm + K)]
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    env.reset_device()
    for i in range(Wm):
        env.step_device(acts[i])
    barrier()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(K)]
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = env.engine.launch_count
    barrier()
    t_wall = time.perf_counter()
    for i in range(K):
        flush.zero_()  # L2 flush, outside the timed event pair
        starts[i].record(stream)
        env.step_device(acts[Wm + i])
        ends[i].record(stream)
    barrier()
    t_wall = time.perf_counter() - t_wall
    launches = env.engine.launch_count - launches0
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    clocks = sampler.stop() if rank == 0 else None
    rew_mean = float(env._reward.mean().item())
    cnt = torch.zeros((B, 8), dtype=torch.int32, device=dev)
    env.engine.get_diag(counters=cnt)
    torch.cuda.synchronize(dev)
    cm = cnt.float().mean(0).tolist()
    solver = {"contacts_last_substep": cm[0], "rows_last_substep": cm[1], "newton_iters_per_step": cm[4],
              "linesearch_iters_per_step": cm[5], "dense_fallback_substeps_per_step": cm[6],
              "lowrank_cross_contacts_per_step": cm[7], "dropped_contacts": cm[3]}

    # ---- end to end through the C-ABI with host buffers (pp_step_host)
    # host buffers are page-locked (the contract's "pinned host memory"): the library DMAs to/from them
    keep = []

    def pinned(shape, dtype):
        t = torch.zeros(shape, dtype=dtype).pin_memory()
        keep.append(t)
        return t.numpy()

    # the same control steps of the episode as the device-timed region: reset, Wm untimed host steps, K timed
    h_act = []
    for i in range(Wm + K):
        buf = pinned((B, 45), torch.float32)
        buf[...] = table[i % period]
        h_act.append(buf)
    h_obs = pinned((B, env.obs_dim), torch.float32)
    h_rew = pinned((B,), torch.float32); h_done = pinned((B,), torch.uint8)
    h_terms = pinned((B, 5), torch.float32); h_disc = pinned((B,), torch.float32)
    env.reset_device()
    for i in range(Wm):
        env.engine.step_host(h_act[i], h_obs, h_rew, h_done, h_terms, h_disc)
    barrier()
    t0 = time.perf_counter()
    for i in range(K):
        env.engine.step_host(h_act[Wm + i], h_obs, h_rew, h_done, h_terms, h_disc)
    barrier()
    e2e_s = time.perf_counter() - t0
    h2d = B * 45 * 4
    d2h = B * (env.obs_dim + 1 + 5 + 1) * 4 + B

    dev_ms, e2e_ms = sharding.max_over_ranks([dev_ms, e2e_s * 1e3], device=dev)
    if rank == 0:
        peaks_path = ROOT / "MEASURED_PEAKS.json"
        if peaks_path.exists():
            peak, peak_src = float(json.loads(peaks_path.read_text())["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        bpe = bytes_per_env_step(env.obs_dim)
        # from the committed `ncu --set full` capture of this kernel on this workload (profiles/): DRAM bytes of one
        # launch, and the figures that say more than the HBM fraction about a latency-bound kernel
        traffic, util = None, {}
        tpath = ROOT / "profiles" / "ncu_bench_kernel.json"
        if tpath.exists() and B == 4096:
            tj = json.loads(tpath.read_text())
            traffic = (tj["dram_bytes_read"] + tj["dram_bytes_write"]) / 1e9  # GB per launch, like achieved * kernel_s
            util = {k: tj[k] for k in ("issue_slot_busy", "lanes_active", "achieved_occupancy", "capture") if k in tj}
        kern_ms = dev_ms / K  # one kernel launch per step: the step IS the dominant kernel
        achieved = bpe * B / (kern_ms * 1e-3) / 1e9
        line = {"metric": METRIC, "value": sharding.whole_job_throughput(shard, K, dev_ms), "unit": UNIT, "n_gpus": world,
                "steps": K, "warmup": Wm, "ms_per_step": dev_ms / K, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": workload_config(B, world, task),
                "e2e": {"value": B * world * K / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
                        "d2h_bytes_per_step": d2h},
                "gpu_launches": int(launches), "clocks": clocks,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "traffic_unit": "GB per launch (ncu)",
                             "algorithmic_gb_per_launch": bpe * B / 1e9, "peak_source": peak_src,
                             "bytes_per_env_step": bpe, "kernel": "pp_hands_step_kernel",
                             "kernel_ms": kern_ms, **util},
                "wall_ms_per_step_incl_flush": 1e3 * t_wall / K, "mean_reward_last_step": rew_mean, "solver": solver}
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            n_envs, steps = 8 * cores, args.cpu_steps
            v = time_cpu_port(task, table, n_envs, steps, 1)
            v_dense = time_cpu_port(task, table, 2 * cores, max(2, steps // 4), 1, sparse=False)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": f"{n_envs} envs x {steps} control steps of the same workload, fp64 oracle "
                                              f"port po_step_sparse (block-structured, no allocation in the step), OpenMP "
                                              f"over envs; MuJoCo unavailable on this box",
                                    "dense_port_value": v_dense,
                                    "dense_port_note": "the clarity-first dense 140x140 po_step (round 1's baseline), "
                                                       "same results to 1e-12"}
    env.close()
    del env, acts, flush
    torch.cuda.empty_cache()
    # ---- the other BASELINE configs, device-resident, each with its own CUDA-event timing (all ranks take part)
    extra, agent = {}, None
    sys.path.insert(0, str(ROOT / "tools"))
    wanted = [c for c in args.extras.split(",") if c]
    if "3" in wanted:
        import bench_configs
        extra["config3"] = bench_configs.measure(3, args.envs3, args.extra_steps, 20, local, world, rank)
    if "5" in wanted:
        import bench_configs
        extra["config5"] = bench_configs.measure(5, args.envs5, args.extra_steps, 20, local, world, rank)
    if "4" in wanted:
        import bench_ppo
        pg = dist.group.WORLD if world > 1 else None
        extra["config4_ppo"], agent = bench_ppo.measure(args.envs, args.extra_steps, 18, local, world, rank, pg)
    if rank == 0:
        if extra:
            line["extra"] = extra
        print(json.dumps(line), flush=True)
    if world > 1:
        if agent is not None:
            # destroy_process_group() blocks after CUDA graphs that captured NCCL kernels were replayed on its
            # communicators: release the graphs, meet at a barrier and leave without it
            agent.close()
            dist.barrier()
            torch.cuda.synchronize()
            sys.stdout.flush()
            os._exit(0)
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="product", choices=["product", "reference"])
    ap.add_argument("--envs", type=int, default=4096, help="envs per GPU")
    ap.add_argument("--cpu-steps", type=int, default=16, help="control steps of the bounded CPU sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--extras", default="3,5,4", help="other BASELINE configs timed after the headline ('' = none)")
    ap.add_argument("--extra-steps", type=int, default=10, help="timed control steps of each extra config")
    ap.add_argument("--envs3", type=int, default=16384, help="envs per GPU of config 3")
    ap.add_argument("--envs5", type=int, default=65536, help="envs per GPU of config 5")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
