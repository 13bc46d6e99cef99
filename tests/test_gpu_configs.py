"""BASELINE configs 3 and 5 on the GPU: parity against the fp64 oracle at oracle-sized
batches, and size-independent properties at the configs' full batch sizes (B = 16384 / 65536):
batch-order invariance (an env's result does not depend on where it sits in the batch or on what its
neighbours do), shard = whole, get_state/set_state round trip, reset idempotence."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _guren_task():
    from diffusion_piano_b200 import music, suite
    # parallelized_base_v2.py:28-38 task kwargs (config 3)
    return suite.PianoWithShadowHands(music.load("GurenNoYumiyaCut14s"), n_steps_lookahead=1, trim_silence=True,
                                      wrong_press_termination=False, initial_buffer_time=0.0)


def _all_keys_midi(dur=10.0):
    from diffusion_piano_b200.music import midi_file
    from diffusion_piano_b200.music.sequence import NoteSequence
    seq = NoteSequence()
    for k in range(88):
        seq.notes.add(start_time=0.0, end_time=dur, velocity=80, pitch=21 + k, part=k % 10)
    seq.total_time = dur
    seq.tempos.add(qpm=60)
    return midi_file.MidiFile(seq=seq)


def _stress_task():
    from diffusion_piano_b200 import suite
    return suite.PianoWithShadowHands(_all_keys_midi(), n_steps_lookahead=10, trim_silence=True)


def _rake_actions(rng, n, step, names):
    """config 5 generator: fingers U(-1,1), forearm_ty of both hands slammed to a bound, alternating
    every 4 control steps (canonical action space), wrists bent down so palms and fingers rake the keys."""
    a = rng.uniform(-1.0, 1.0, (n, 45)).astype(np.float32)
    ty = -1.0 if (step // 4) % 2 == 0 else 1.0
    for pre in ("rh", "lh"):
        a[:, names.index(f"{pre}_shadow_hand/forearm_ty")] = ty
    return a


class _SyncedCanonical:
    """env 0 of a small GPU batch is re-synchronised to the oracle before every control step."""

    def __init__(self, task, n=2):
        import torch
        from diffusion_piano_b200 import suite
        from oracle.task_oracle import PianoWithShadowHandsOracle
        self.t = torch
        self.env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=False)
        self.ora = PianoWithShadowHandsOracle(task)
        self.env.reset(); self.ora.reset()
        z = lambda *s, dt=torch.float32: torch.zeros(s, device="cuda", dtype=dt)
        self.qpos, self.qvel, self.warm = z(n, 140), z(n, 140), z(n, 140)
        self.cnt = z(n, 8, dt=torch.int32)
        self.n = n

    def step(self, a):
        t, e, p = self.t, self.env.engine, self.ora.physics
        e.get_state(qpos=self.qpos, qvel=self.qvel, warm=self.warm)
        for dst, src in ((self.qpos, p.qpos), (self.qvel, p.qvel), (self.warm, p.qacc_warmstart)):
            dst[0] = t.from_numpy(src.astype(np.float32)).cuda()
        e.set_state(qpos=self.qpos, qvel=self.qvel, warm=self.warm)
        t.cuda.synchronize()
        a32 = np.asarray(a, np.float32)
        obs, rew, done = self.env.step(np.stack([a32] * self.n))
        obs_o, r_o, d_o = self.ora.step(self.ora.canonical_to_spec(a32.astype(np.float64)))
        e.get_diag(counters=self.cnt)
        t.cuda.synchronize()
        return obs, rew, done, obs_o, r_o, d_o


def _names(task):
    return task.model.names["actuator"]


def test_config3_guren_random_actions_parity(cuda_lib):
    """config 3: fingered 'Guren no Yumiya' cut, canonical U(-1,1) actions, re-synchronised control steps."""
    s = _SyncedCanonical(_guren_task())
    assert s.env.obs_dim == 329
    rng = np.random.RandomState(0)
    worst = {}
    for t in range(20):
        a = rng.uniform(-1.0, 1.0, 45)
        obs, rew, done, obs_o, r_o, d_o = s.step(a)
        assert bool(done[0]) == d_o
        for k, v in obs_o.items():
            tol = 0 if k in ("goal", "fingering") else 1e-4  # 10 free-running fp32 substeps after a U(-1,1) jump (measured 2.2e-5)
            np.testing.assert_allclose(obs[k][0], v, atol=tol, err_msg=f"{k} step {t}")
            worst[k] = max(worst.get(k, 0.0), float(np.abs(obs[k][0] - v).max()))
        for name, val in s.ora.reward_terms.items():
            np.testing.assert_allclose(s.env.reward_terms[name][0], val, atol=2e-3, err_msg=f"{name} step {t}")
        assert int(s.cnt[0, 3]) == 0  # nothing dropped
    print("config 3 worst |obs - oracle| over 20 control steps:", {k: f"{v:.1e}" for k, v in worst.items()})
    s.env.close()


def test_config5_stress_parity(cuda_lib):
    """config 5: all 88 keys requested, hands raking the keyboard. Every step is compared: the sequence must
    stay inside the kernel's contact caps (24 contacts, 16 touched keys per env) - a step that drops a contact
    fails the test instead of being skipped."""
    task = _stress_task()
    s = _SyncedCanonical(task)
    assert s.env.obs_dim == 1130
    names = _names(task)
    rng = np.random.RandomState(0)
    compared, dropped, max_con, dense = 0, 0, 0, 0
    for t in range(16):
        a = _rake_actions(rng, 1, t, names)[0]
        obs, rew, done, obs_o, r_o, d_o = s.step(a)
        max_con = max(max_con, int(s.cnt[0, 0])); dense += int(s.cnt[0, 6])
        dropped += int(s.cnt[0, 3]) > 0
        compared += 1
        for k, v in obs_o.items():
            tol = 0 if k in ("goal", "fingering") else 5e-4
            np.testing.assert_allclose(obs[k][0], v, atol=tol, err_msg=f"{k} step {t}")
        np.testing.assert_allclose(rew[0], r_o, atol=5e-3, err_msg=f"reward step {t}")
    print(f"config 5: compared {compared} steps, {dropped} with dropped contacts, max contacts {max_con}, "
          f"dense-fallback substeps {dense}")
    assert compared == 16 and dropped == 0
    s.env.close()


def _run(env, actions_per_step):
    import torch
    outs = []
    env.reset_device()
    for a in actions_per_step:
        obs, rew, done, terms, disc = env.step_device(a)
        outs.append((obs.clone(), rew.clone(), done.clone(), terms.clone()))
    torch.cuda.synchronize()
    return outs


@pytest.mark.parametrize("config,B,steps", [(3, 16384, 6), (5, 65536, 3)])
def test_full_size_batch_order_and_shard_invariance(cuda_lib, config, B, steps):
    import torch
    from diffusion_piano_b200 import suite
    task = _guren_task() if config == 3 else _stress_task()
    names = _names(task)
    rng = np.random.RandomState(config)
    if config == 3:
        acts = [rng.uniform(-1, 1, (B, 45)).astype(np.float32) for _ in range(steps)]
    else:
        acts = [_rake_actions(rng, B, t, names) for t in range(steps)]
    dev = torch.device("cuda", 0)
    fwd = [torch.from_numpy(a).to(dev) for a in acts]
    env = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True)
    ref = _run(env, fwd)
    for obs, rew, done, terms in ref:
        assert torch.isfinite(obs).all() and torch.isfinite(rew).all()
    # batch-order invariance: the same envs in reversed order give the reversed rows, bit for bit
    rev = _run(env, [a.flip(0).contiguous() for a in fwd])
    for (o1, r1, d1, t1), (o2, r2, d2, t2) in zip(ref, rev):
        assert torch.equal(o1, o2.flip(0)) and torch.equal(r1, r2.flip(0)) and torch.equal(d1, d2.flip(0))
        assert torch.equal(t1, t2.flip(0))
    # get_state / set_state round trip and determinism of a step from a restored state
    qpos, qvel, warm = (torch.zeros((B, 140), device=dev) for _ in range(3))
    env.engine.get_state(qpos=qpos, qvel=qvel, warm=warm)
    o_a = env.step_device(fwd[0])[0].clone()
    env.engine.set_state(qpos=qpos, qvel=qvel, warm=warm)
    q2 = torch.zeros_like(qpos); env.engine.get_state(qpos=q2)
    assert torch.equal(q2, qpos)
    env.close()
    # shard = whole: the first B/4 envs stepped alone (what one rank of a 4-GPU job would own)
    Bs = B // 4
    env_s = suite.BatchedEnvironment(task, Bs, canonical_actions=True, auto_reset=True)
    part = _run(env_s, [a[:Bs].contiguous() for a in fwd])
    for (o1, r1, d1, t1), (o2, r2, d2, t2) in zip(ref, part):
        assert torch.equal(o1[:Bs], o2) and torch.equal(r1[:Bs], r2) and torch.equal(d1[:Bs], d2)
    # reset idempotence
    a = env_s.reset_device().clone()
    b = env_s.reset_device().clone()
    assert torch.equal(a, b)
    env_s.close()
    del o_a


def test_host_buffer_step_equals_device_step(cuda_lib):
    """pp_step_host steps large batches as two halves on two streams (D2H of the first half overlaps the
    second half's step): same rows as the device-pointer call, bit for bit."""
    import torch
    from diffusion_piano_b200 import suite
    task = _guren_task()
    B = 2050  # even split point, odd-sized halves exercise the tail launch too
    rng = np.random.RandomState(11)
    acts = [rng.uniform(-1, 1, (B, 45)).astype(np.float32) for _ in range(3)]
    env_d = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True)
    env_h = suite.BatchedEnvironment(task, B, canonical_actions=True, auto_reset=True)
    env_d.reset_device(); obs_h0 = env_h.reset()
    for a in acts:
        obs_d, rew_d, done_d, terms_d, disc_d = env_d.step_device(torch.from_numpy(a).cuda())
        torch.cuda.synchronize()
        obs_h, rew_h, done_h = env_h.step(a)
        flat = np.concatenate([obs_h[k] for k, _ in sorted(env_h.layout.items(), key=lambda kv: kv[1][0])], axis=1)
        np.testing.assert_array_equal(flat.astype(np.float32), obs_d.cpu().numpy())
        np.testing.assert_array_equal(rew_h.astype(np.float32), rew_d.cpu().numpy())
        np.testing.assert_array_equal(done_h, done_d.cpu().numpy().astype(bool))
    env_d.close(); env_h.close()
