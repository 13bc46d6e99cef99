"""The launch shape of the two-hands step (pp_set_schedule: envs per block, work-ordered envs, the auto-tuner)
never changes a result: an env's step depends on its own state and action only."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _run(schedule, n=333, steps=20):
    import torch

    from diffusion_piano_b200 import music, suite

    task = suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True)
    env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, schedule=schedule)
    g = torch.Generator(device="cuda")
    g.manual_seed(1)
    env.reset_device()
    out = []
    for _ in range(steps):
        a = torch.rand((n, 45), device="cuda", generator=g) * 2 - 1
        obs, rew, done, terms, _ = env.step_device(a)
        out.append((obs.cpu().numpy().copy(), rew.cpu().numpy().copy(), terms.cpu().numpy().copy()))
    sched = env.engine.get_schedule()
    cnt = torch.zeros((n, 8), dtype=torch.int32, device="cuda")
    env.engine.get_diag(counters=cnt)
    env.close()
    return out, sched, cnt.cpu().numpy()


def test_schedules_give_identical_results(monkeypatch):
    monkeypatch.delenv("PP_SCHEDULE", raising=False)  # the suite-wide knob (pp_create) would replace "auto"
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    ref, s0, c0 = _run("lockstep")
    assert s0 == {"mode": "lockstep", "block_envs": 2, "sorted_by_work": 0, "in_force": "lockstep"}
    div, s1, c1 = _run("divergent")
    assert s1 == {"mode": "divergent", "block_envs": 10, "sorted_by_work": 1, "in_force": "divergent"}
    auto, s2, _ = _run("auto")  # 20 steps: tuned after step 16, either shape may win at this size
    assert s2["mode"] == "auto" and s2["block_envs"] in (2, 10)
    # per substep a prologue / Newton / finish launch, the Newton warps pulling envs from cost-class lists: the same
    # code with the other phases compiled out, hence the same bits
    comp, s3, c3 = _run("compacted")
    assert s3["mode"] == "compacted" == s3["in_force"]
    np.testing.assert_array_equal(c0, c3)
    for other in (div, auto, comp):
        for (o0, r0, t0), (o1, r1, t1) in zip(ref, other):
            np.testing.assert_array_equal(o0, o1)
            np.testing.assert_array_equal(r0, r1)
            np.testing.assert_array_equal(t0, t1)
    np.testing.assert_array_equal(c0, c1)  # same Newton / line-search iteration counts per env
    assert c0[:, 4].std() > 0  # the envs really did different amounts of work


def test_compacted_schedule_with_resets_and_options():
    """The compacted schedule through episode ends (auto-reset at substep 0 of the next step), the optimal-transport
    reward, the one-substep-old derived quantities of legacy_step=False and a batch that is no multiple of 10."""
    import torch
    from diffusion_piano_b200 import music, suite

    task = suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), n_steps_lookahead=2, control_timestep=0.25,
                                      disable_fingering_reward=True, wrong_press_termination=True)
    outs = {}
    for sched in ("lockstep", "compacted"):
        env = suite.BatchedEnvironment(task, 57, canonical_actions=True, auto_reset=True, schedule=sched,
                                       legacy_step=False)
        g = torch.Generator(device="cuda")
        g.manual_seed(7)
        env.reset_device()
        rows, n_done = [], 0
        for _ in range(30):
            a = torch.rand((57, 45), device="cuda", generator=g) * 2 - 1
            obs, rew, done, terms, disc = env.step_device(a)
            n_done += int(done.sum())
            rows.append([x.cpu().numpy().copy() for x in (obs, rew, done, terms, disc)])
        outs[sched] = rows
        env.close()
        assert n_done > 57  # every env ended at least one episode (wrong presses end them early)
    for a, b in zip(outs["lockstep"], outs["compacted"]):
        for x, y in zip(a, b):
            np.testing.assert_array_equal(x, y)


@pytest.mark.parametrize("n", [1, 9, 10, 11, 148 * 10 + 1, 1999])
def test_compacted_schedule_batch_sizes(n, monkeypatch):
    """Batch sizes around the block shape (10 envs) and the persistent grid (one block per SM): spare warps in the last
    block, fewer envs than warps, more blocks than SMs. Two handles step at the same time on different streams (the
    host-buffer API runs on the handle's own non-blocking stream)."""
    import torch
    from diffusion_piano_b200 import music, suite

    monkeypatch.delenv("PP_SCHEDULE", raising=False)
    task = suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), n_steps_lookahead=1)
    ref = suite.BatchedEnvironment(task, n, canonical_actions=True, schedule="lockstep")
    dev = suite.BatchedEnvironment(task, n, canonical_actions=True, schedule="compacted")
    host = suite.BatchedEnvironment(task, n, canonical_actions=True, schedule="compacted")
    ref.reset_device(); dev.reset_device(); host.reset()
    rng = np.random.RandomState(n)
    for _ in range(4):
        a = rng.uniform(-1, 1, (n, 45)).astype(np.float32)
        at = torch.as_tensor(a, device="cuda")
        o_d, r_d, d_d, t_d, _ = dev.step_device(at)  # asynchronous: still running while `host` steps
        o_h, r_h, d_h = host.step(a)
        o_r, r_r, d_r, t_r, _ = ref.step_device(at)
        flat = np.concatenate([o_h[k] for k, _ in sorted(host.layout.items(), key=lambda kv: kv[1][0])], axis=1)
        assert torch.equal(o_d, o_r) and torch.equal(r_d, r_r) and torch.equal(t_d, t_r) and torch.equal(d_d, d_r)
        np.testing.assert_array_equal(flat, o_r.cpu().numpy().astype(np.float64))
        np.testing.assert_array_equal(r_h, r_r.cpu().numpy().astype(np.float64))
        assert np.isfinite(flat).all()
    for e in (ref, dev, host):
        e.close()
