"""Host-side multi-GPU plumbing of the PPO agent (one process per GPU; NCCL on the GPUs, gloo in the CPU tests).

The reference is single-process: its `update` sees the whole batch of a control step (`ppo_v2.py:211-249`). With
env shards, the per-env vectors the batch preparation needs - raw reward, done, value, next value: 4 floats per
env - are all-gathered so that every rank prepares the GLOBAL batch (rank-major env order = the unsharded order)
and takes its own slice of returns / advantages; gradients are then summed over ranks and scaled by 1 / world
inside the clip + Adam kernel."""

from __future__ import annotations

import torch


def gather_global_batch(packed: torch.Tensor, gathered: torch.Tensor, glob: torch.Tensor, group) -> torch.Tensor:
    """packed [4, n] (this rank) -> glob [4, world * n], row r = concatenation over ranks of packed[r].
    `gathered` [world, 4, n] and `glob` are caller-owned work buffers (no allocation: graph / stream friendly)."""
    world, four, n = gathered.shape
    torch.distributed.all_gather_into_tensor(gathered.view(-1), packed.reshape(-1), group=group)
    glob.view(four, world, n).copy_(gathered.permute(1, 0, 2))
    return glob


def shard_slice(rank: int, n: int) -> slice:
    """Rows of the global batch owned by `rank` (n envs per rank)."""
    return slice(rank * n, (rank + 1) * n)


def clip_coefficient(summed_grad_norm2: float, world: int, max_norm: float) -> float:
    """Factor applied to the SUMMED gradient: 1 / world (the mean over ranks) times clip_grad_norm_'s coefficient
    for the mean gradient, whose norm is sqrt(norm2) / world. Mirrors `clip_adam_kernel` (csrc/ppo_kernels.cu)."""
    scale = 1.0 / world
    total = summed_grad_norm2 ** 0.5 * scale
    return min(max_norm / (total + 1e-6), 1.0) * scale


def sync_scheduler_metric(ctrl: torch.Tensor, byte_offset: int, world: int, group) -> None:
    """Mean over ranks, in place, of ONE float32 inside the device control block (`ctrl`: uint8 tensor).
    The critic's ReduceLROnPlateau (`ppo_v2.py:305-308`) steps on the critic loss of the LAST minibatch, which each
    rank computes on its own shard with its own dropout draws: without this the ranks' `best` / `num_bad` / `lr`
    drift apart and the critic replicas then apply one averaged gradient with different learning rates. After the
    all-reduce every rank holds the same bits, so the scheduler state stays identical on all ranks."""
    view = ctrl[byte_offset:byte_offset + 4].view(torch.float32)
    torch.distributed.all_reduce(view, group=group)
    view.mul_(1.0 / world)


def all_ranks_agree(flag: bool, device, group) -> bool:
    """True only when EVERY rank passes True (all-reduce MIN of one integer). The training loop's `if all(dones):
    break` (`parallelized_base_v2.py:163`) must be one decision for all ranks: `agent.update` holds collectives,
    and a rank that left the loop alone would leave the others waiting in them."""
    t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=device)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN, group=group)
    return bool(t.item())
