"""N > 1 host logic on CPU: two gloo ranks. The step has no collective; what is
checked is the env partition, the per-rank RNG streams, the max-over-ranks timing / sum-of-statistics
reductions `bench.py` and a training loop use, and that envs stepped shard-by-shard (here with the CPU
oracle standing in for the device step) give exactly the rows of the unsharded batch."""
import os
import socket

import numpy as np
import pytest


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _per_env_actions(global_env: int, step: int) -> np.ndarray:
    """Deterministic per-(env, step) action in the self-actuated piano's spec: a function of the GLOBAL
    env index, so that the result cannot depend on how envs are spread over ranks."""
    rng = np.random.RandomState(1000 * global_env + step)
    a = rng.uniform(0.0, 0.06, 89)
    a[88] = float(rng.uniform() > 0.5)
    return a


def _step_envs(env_ids, n_steps: int) -> np.ndarray:
    """rows [len(env_ids), 2 + 88]: return, done, final key state - from the CPU oracle."""
    from diffusion_piano_b200 import music, suite
    from oracle.task_oracle import SelfActuatedPianoOracle

    task = suite.SelfActuatedPiano(music.load("CMajorScaleOneHand"), n_steps_lookahead=1, control_timestep=0.05)
    rows = []
    for g in env_ids:
        ora = SelfActuatedPianoOracle(task)
        ora.reset()
        ret, done = 0.0, False
        for t in range(n_steps):
            _, r, done = ora.step(_per_env_actions(g, t))
            ret += r
        rows.append(np.concatenate([[ret, float(done)], ora.physics.qpos[:88]]))
    return np.asarray(rows)


def _worker(rank: int, world: int, port: int, envs_per_rank: int, n_steps: int, q) -> None:
    import torch.distributed as dist

    from diffusion_piano_b200 import sharding

    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = sharding.shard_from_env(envs_per_rank, seed=7)
        local = _step_envs(sh.env_range, n_steps)
        rows = sharding.gather_rows(local)
        # device-timed milliseconds would go here; every rank must see the slowest rank's time
        tmax = sharding.max_over_ranks([10.0 + rank, 3.0 - rank])
        stats = sharding.sum_over_ranks([local[:, 0].sum(), len(local)])
        draw = float(sh.random_state().uniform())
        draws = [None] * world
        dist.all_gather_object(draws, draw)
        q.put((rank, list(sh.env_range), sh.seed, rows, tmax, stats, draws,
               sharding.whole_job_throughput(sh, n_steps, tmax[0])))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_env_sharding_gloo():
    import torch.multiprocessing as mp

    world, B, n_steps = 2, 3, 4
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, B, n_steps, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted((q.get(timeout=240) for _ in range(world)), key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    whole = _step_envs(range(world * B), n_steps)  # the unsharded batch
    seen = []
    for rank, env_range, seed, rows, tmax, stats, draws, thr in out:
        assert env_range == list(range(rank * B, (rank + 1) * B))
        assert seed == 7 + rank
        seen += env_range
        np.testing.assert_array_equal(rows, whole)  # bit-exact: no data-path exchange, same arithmetic
        assert tmax == [10.0 + world - 1, 3.0]
        np.testing.assert_allclose(stats, [whole[:, 0].sum(), world * B], rtol=1e-12)
        assert len(set(draws)) == world  # per-rank RNG streams differ
        assert thr == pytest.approx(world * B * n_steps / (tmax[0] * 1e-3))
    assert sorted(seen) == list(range(world * B))


def test_single_process_helpers_are_identity():
    from diffusion_piano_b200 import sharding

    sh = sharding.shard_from_env(8, seed=3, environ={})
    assert (sh.rank, sh.world_size, sh.global_envs, list(sh.env_range)) == (0, 1, 8, list(range(8)))
    assert sharding.max_over_ranks([1.5, 2.5]) == [1.5, 2.5]
    assert sharding.sum_over_ranks([4.0]) == [4.0]
    with pytest.raises(ValueError):
        sharding.shard_from_env(8, environ={"RANK": "2", "WORLD_SIZE": "2"})
