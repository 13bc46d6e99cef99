"""Multi-GPU layout of the batched step: environments are independent, so rank r of W owns the
global env indices [r*B, (r+1)*B) with its own `pp_env` handle and RNG stream `seed + r`
(SURVEY.md §8e). There is NO collective inside the step; torch.distributed is only used for what
a caller does around it: the max-over-ranks timing of `bench.py` and the sums of episode
statistics / gradient buckets of a training loop (reference parallelized_base_v2.py:116-166 runs
those on one process).

Everything here works on any backend (NCCL on the GPU box, gloo in the CPU tests).
"""

from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, Optional, Sequence

import numpy as np


@dataclass(frozen=True)
class Shard:
    rank: int
    world_size: int
    local_rank: int
    envs_per_rank: int
    seed: int

    @property
    def global_envs(self) -> int:
        return self.envs_per_rank * self.world_size

    @property
    def env_range(self) -> range:
        """Global indices of the envs this rank owns."""
        return range(self.rank * self.envs_per_rank, (self.rank + 1) * self.envs_per_rank)

    def random_state(self) -> np.random.RandomState:
        return np.random.RandomState(self.seed)


def shard_from_env(envs_per_rank: int, seed: int = 0, environ: Optional[Dict[str, str]] = None) -> Shard:
    """Reads RANK / WORLD_SIZE / LOCAL_RANK as torchrun sets them (1 process per GPU)."""
    e = os.environ if environ is None else environ
    rank, world = int(e.get("RANK", "0")), int(e.get("WORLD_SIZE", "1"))
    local = int(e.get("LOCAL_RANK", str(rank)))
    if not 0 <= rank < world:
        raise ValueError(f"RANK {rank} outside WORLD_SIZE {world}")
    if envs_per_rank <= 0:
        raise ValueError("envs_per_rank must be positive")
    return Shard(rank, world, local, int(envs_per_rank), int(seed) + rank)


def _dist():
    import torch.distributed as dist

    return dist


def _active() -> bool:
    dist = _dist()
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def max_over_ranks(values: Sequence[float], device=None) -> list:
    """Element-wise MAX over ranks (device-timed milliseconds of a step: the job is as slow as
    its slowest rank). Identity on one process."""
    import torch

    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if _active():
        _dist().all_reduce(t, op=_dist().ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


def sum_over_ranks(values: Sequence[float], device=None) -> list:
    """Element-wise SUM over ranks (episode statistics: sum of returns, of reward terms, counts)."""
    import torch

    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if _active():
        _dist().all_reduce(t, op=_dist().ReduceOp.SUM)
    return [float(x) for x in t.tolist()]


def whole_job_throughput(shard: Shard, steps: int, ms_max_over_ranks: float) -> float:
    """env-steps/s of the whole job: every rank stepped `envs_per_rank` envs `steps` times."""
    return shard.global_envs * steps / (ms_max_over_ranks * 1e-3)


def gather_rows(local: np.ndarray) -> np.ndarray:
    """All ranks' per-env rows concatenated in global env order (host arrays; used by tests and
    evaluation code, never by the step)."""
    if not _active():
        return np.asarray(local)
    out = [None] * _dist().get_world_size()
    _dist().all_gather_object(out, np.asarray(local))
    return np.concatenate(out, axis=0)
