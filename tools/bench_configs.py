"""Throughput of the BASELINE.json configs that `bench.py` does not time (it times configs[1]):
  1  self-actuated piano, TwinkleTwinkleRousseau, goal-following policy            (B = --envs1, default 4096)
  3  'Guren no Yumiya Cut 14s' + fingering, U(-1,1) canonical actions, L = 1       (B = 16384)
  5  contact-heavy stress: all 88 keys, L = 10, raking hands                       (B = 65536 per GPU)
One JSON line per config (rank 0), device-resident, CUDA events, max over ranks; the launch shape is the handle's
auto-tuner (warm-up >= 18 steps so that it has finished). `python tools/bench_configs.py [--configs 1,3,5]`;
under torchrun every rank owns the full per-GPU batch (weak scaling, no collective in the step)."""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def all_keys_midi(dur=10.0):
    from diffusion_piano_b200.music import midi_file
    from diffusion_piano_b200.music.sequence import NoteSequence

    seq = NoteSequence()
    for k in range(88):
        seq.notes.add(start_time=0.0, end_time=dur, velocity=80, pitch=21 + k, part=k % 10)
    seq.total_time = dur
    seq.tempos.add(qpm=60)
    return midi_file.MidiFile(seq=seq)


def measure(cfg, B, steps, warmup, local, world, rank, schedule="auto"):
    """One config, device-resident, CUDA events over `steps` steps after `warmup`, max over ranks -> JSON-able dict."""
    import torch

    from diffusion_piano_b200 import music, suite

    dev = torch.device("cuda", local)
    g = torch.Generator(device=dev)
    g.manual_seed(1234 + rank)
    if cfg == 1:
        task = suite.SelfActuatedPiano(music.load("TwinkleTwinkleRousseau"), control_timestep=0.01)
        env = suite.BatchedEnvironment(task, B, device=local, auto_reset=True)
        name = "self-actuated piano, TwinkleTwinkleRousseau, goal-following policy"
    elif cfg == 3:
        task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)
        env = suite.BatchedEnvironment(task, B, device=local, canonical_actions=True, auto_reset=True, schedule=schedule)
        name = "PianoWithShadowHands 'Guren no Yumiya Cut 14s' + fingering, U(-1,1) actions"
    else:
        task = suite.PianoWithShadowHands(all_keys_midi(), n_steps_lookahead=10, trim_silence=True)
        env = suite.BatchedEnvironment(task, B, device=local, canonical_actions=True, auto_reset=True, schedule=schedule)
        name = "contact-heavy stress: all 88 keys, L=10, forearm_ty slammed to alternating bounds, U(-1,1) fingers"
    ad = env.action_dim
    obs = env.reset_device()
    hi = torch.as_tensor(env.engine.action_maximum, device=dev)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = 0
    for it in range(warmup + steps):
        if cfg == 1:  # goal-following: press what the goal row asks (examples/self_actuated_piano_env.py:82-96)
            goal = obs[:, :89]
            a = torch.where(goal > 0, hi.expand(B, -1), torch.zeros((), device=dev)).contiguous()
        else:
            a = torch.rand((B, ad), device=dev, generator=g) * 2 - 1
            if cfg == 5:
                a[:, 21] = a[:, 43] = -1.0 if (it // 4) % 2 == 0 else 1.0
        if it == warmup:
            if world > 1:
                torch.distributed.barrier()
            torch.cuda.synchronize()
            launches0 = env.engine.launch_count
            ev0.record()
        obs = env.step_device(a)[0]
    ev1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([ev0.elapsed_time(ev1)], device=dev)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    ms = float(ms.item())
    line = {"config": cfg, "metric": "env steps/sec", "value": B * world * steps / (ms * 1e-3), "unit": "env-steps/s",
            "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": ms / steps,
            "envs_per_gpu": B, "obs_dim": env.obs_dim, "workload": name, "data": "synthetic",
            "gpu_launches": env.engine.launch_count - launches0,
            "note": "timed region includes the on-device action generation (torch.rand / torch.where)"}
    if cfg != 1:
        line["schedule"] = env.engine.get_schedule()
        cnt = torch.zeros((B, 8), dtype=torch.int32, device=dev)
        env.engine.get_diag(counters=cnt)
        c = cnt.float()
        m = c.mean(0).cpu().numpy()
        line["solver"] = {"contacts_last_substep": float(m[0]), "newton_iters_per_step": float(m[4]),
                          "dense_fallback_substeps_per_step": float(m[6]), "dropped_contacts": float(m[3]),
                          "envs_with_dropped_contacts": float((c[:, 3] > 0).float().mean().item()),
                          "contacts_max": int(c[:, 0].max().item()),
                          "contacts_histogram_0_2_4_8_16_24": torch.histc(c[:, 0], bins=5, min=0, max=24).int().tolist()}
    env.close()
    del env
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="1,3,5")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--envs1", type=int, default=4096)
    ap.add_argument("--envs3", type=int, default=16384)
    ap.add_argument("--envs5", type=int, default=65536)
    ap.add_argument("--schedule", default="auto")
    args = ap.parse_args()
    import torch

    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
    for cfg in (int(c) for c in args.configs.split(",")):
        B = {1: args.envs1, 3: args.envs3}.get(cfg, args.envs5)
        line = measure(cfg, B, args.steps, args.warmup, local, world, rank, args.schedule)
        if rank == 0:
            print(json.dumps(line), flush=True)
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
