"""PPO with env shards, host logic on CPU with two gloo ranks: the all-gather of the per-env vectors
reproduces the unsharded batch, so the batch preparation (normaliser, index-wise recursion, advantage statistics;
here the oracle's) gives every rank exactly its rows of the single-process result; and the factor the clip + Adam
kernel applies to the summed gradient equals clip_grad_norm_ on the mean gradient."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _global_vectors(n_global: int) -> np.ndarray:
    rs = np.random.RandomState(5)
    out = rs.randn(4, n_global).astype(np.float32)
    out[1] = (rs.rand(n_global) < 0.1).astype(np.float32)  # dones
    return out


def _worker(rank: int, world: int, port: int, n: int, q):
    import torch
    import torch.distributed as td

    from diffusion_piano_b200.ppo import dist
    from oracle import ppo_oracle as po

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        whole = _global_vectors(n * world)
        packed = torch.from_numpy(whole[:, rank * n:(rank + 1) * n].copy())
        gathered, glob = torch.empty(world, 4, n), torch.empty(4, world * n)
        g = dist.gather_global_batch(packed, gathered, glob, td.group.WORLD).numpy()
        rms = po.RunningMeanStd()
        rn = rms(g[0].astype(np.float64)).astype(np.float32).astype(np.float64)
        ret, _, adv = po.prepare_batch(rn, g[1].astype(np.float64), g[2].astype(np.float64), g[3].astype(np.float64), 0.99)
        mine = dist.shard_slice(rank, n)
        # gradient averaging: sum over ranks, then the kernel's factor
        grad = torch.full((7,), float(rank + 1))
        td.all_reduce(grad)
        coef = dist.clip_coefficient(float((grad * grad).sum()), world, 0.5)
        q.put((rank, g.copy(), ret[mine].copy(), adv[mine].copy(), (grad * coef).numpy().copy()))
    finally:
        td.destroy_process_group()


def test_two_rank_batch_preparation_equals_single_process():
    import torch.multiprocessing as mp

    from oracle import ppo_oracle as po

    world, n = 2, 37
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = sorted((q.get(timeout=120) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole = _global_vectors(n * world)
    rms = po.RunningMeanStd()
    rn = rms(whole[0].astype(np.float64)).astype(np.float32).astype(np.float64)
    ret, _, adv = po.prepare_batch(rn, whole[1].astype(np.float64), whole[2].astype(np.float64),
                                   whole[3].astype(np.float64), 0.99)
    mean_grad = {"g": np.full(7, 1.5)}  # mean of rank gradients 1 and 2
    po.clip_grad_norm(mean_grad, 0.5)
    for rank, g, ret_r, adv_r, clipped in results:
        np.testing.assert_array_equal(g, whole)  # rank-major order = the unsharded order
        np.testing.assert_array_equal(ret_r, ret[rank * n:(rank + 1) * n])
        np.testing.assert_array_equal(adv_r, adv[rank * n:(rank + 1) * n])
        np.testing.assert_allclose(clipped, mean_grad["g"], rtol=1e-6)


def test_clip_coefficient_single_rank_is_clip_grad_norm():
    from diffusion_piano_b200.ppo import dist
    from oracle import ppo_oracle as po

    for norm in (0.1, 0.5, 3.0):
        g = {"g": np.array([norm])}
        po.clip_grad_norm(g, 0.5)
        assert abs(dist.clip_coefficient(norm * norm, 1, 0.5) * norm - g["g"][0]) < 1e-12


def _plateau(state, metric, patience=100, factor=0.5):
    """plateau_kernel of csrc/ppo_kernels.cu for the critic ('min' mode, relative threshold 1e-4), in float32."""
    lr, best, bad = state
    metric = np.float32(metric)
    if metric < best * np.float32(1.0 - 1e-4):
        best, bad = metric, 0
    else:
        bad += 1
    if bad > patience:
        nlr = np.float32(lr * np.float32(factor))
        if lr - nlr > 1e-8:
            lr = nlr
        bad = 0
    return lr, best, bad


def _sched_worker(rank: int, world: int, port: int, q):
    import ctypes

    import torch
    import torch.distributed as td

    from diffusion_piano_b200.ppo import _kernels, dist

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    td.init_process_group("gloo", rank=rank, world_size=world)
    try:
        off = _kernels.PPOCtrl.critic_loss.offset
        ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8)
        loss = ctrl[off:off + 4].view(torch.float32)
        rng = np.random.RandomState(100 + rank)  # every rank's own shard and dropout draws: its own critic loss
        synced = (np.float32(2e-4), np.float32(np.inf), 0)
        local = synced
        for step in range(260):  # more than two patience windows of a loss that stops improving after step 20
            mine = 1.0 + 5.0 / (1 + min(step, 20)) + 0.05 * rng.standard_normal()
            local = _plateau(local, mine)
            loss.fill_(float(mine))
            dist.sync_scheduler_metric(ctrl, off, world, td.group.WORLD)
            synced = _plateau(synced, float(loss.item()))
        # the loop-exit decision: rank 1 alone sees an early termination
        agree_mixed = dist.all_ranks_agree(rank == 1, torch.device("cpu"), td.group.WORLD)
        agree_all = dist.all_ranks_agree(True, torch.device("cpu"), td.group.WORLD)
        q.put((rank, [float(x) for x in synced], [float(x) for x in local], agree_mixed, agree_all,
               bytes(ctrl.numpy().tobytes())))
    finally:
        td.destroy_process_group()


def test_two_rank_scheduler_state_stays_identical():
    """the critic's plateau scheduler steps on a per-rank metric. With the metric averaged over
    ranks first (`dist.sync_scheduler_metric`, used by PPOAgent.update) lr / best / num_bad are bit-identical on
    all ranks after more than `patience` non-improving updates; without it they are not (same draws). Also the
    training loop's global break decision (`dist.all_ranks_agree`)."""
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_sched_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, s0, l0, mixed0, all0, c0), (_, s1, l1, mixed1, all1, c1) = out
    assert s0 == s1 and c0 == c1          # synchronised metric: identical scheduler state and control-block bytes
    assert s0[0] < 2e-4                   # the learning rate did halve at least once (patience 100 exceeded)
    assert l0 != l1                       # the unsynchronised schedulers drifted apart (best differs)
    assert mixed0 is False and mixed1 is False and all0 is True and all1 is True
