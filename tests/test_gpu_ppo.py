"""PPO v2 on the device against the oracle (oracle/ppo_oracle.py, itself pinned against a
recorded run of the reference's ppo_v2.PPOAgent.update) and against the reference's recorded results directly.
Every kernel is called through the C-ABI of libppo_kernels.so."""
import ctypes
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return torch


def _agent(torch, state_dim, action_dim, actor=None, critic=None, **kw):
    from diffusion_piano_b200.ppo import PPOAgent

    ag = PPOAgent(state_dim, action_dim, device="cuda:0", **kw)
    if actor is not None:
        ag.actor.load(actor)
        ag.critic.load(critic)
    return ag


def _stream(torch):
    return torch.cuda.current_stream().cuda_stream


def test_block_forward_backward_matches_oracle(torch_cuda):
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels
    from oracle import ppo_oracle as po

    lib = _kernels.load()
    rs = np.random.RandomState(0)
    for rows, H, p in ((128, 256, 0.1), (44, 128, 0.0), (300, 256, 0.1)):
        y = rs.randn(rows, H).astype(np.float32)
        bias, gamma, beta = (0.1 * rs.randn(H)).astype(np.float32), (1 + 0.1 * rs.randn(H)).astype(np.float32), \
            (0.1 * rs.randn(H)).astype(np.float32)
        mask = (rs.rand(rows, H) >= p).astype(np.uint8)
        dout = rs.randn(rows, H).astype(np.float32)
        d = lambda a: torch.as_tensor(a, device="cuda")  # noqa: E731
        ty, tb, tg, tbe, tm, tdo = d(y), d(bias), d(gamma), d(beta), d(mask), d(dout)
        out, h, stats = torch.empty_like(ty), torch.empty_like(ty), torch.empty(rows, 2, device="cuda")
        mask_out = torch.zeros_like(tm)
        _kernels.check(lib.ppo_relu_ln_fwd(ty.data_ptr(), tb.data_ptr(), tg.data_ptr(), tbe.data_ptr(), out.data_ptr(),
                                           h.data_ptr(), stats.data_ptr(), rows, H, p, tm.data_ptr() if p > 0 else None,
                                           mask_out.data_ptr() if p > 0 else None, None, 0, _stream(torch)))
        # oracle: the block with W = I so that y is the pre-bias Linear output
        o_ref, cache = po._block_forward(y.astype(np.float64), np.eye(H), bias.astype(np.float64),
                                         gamma.astype(np.float64), beta.astype(np.float64))
        if p > 0:
            o_ref = o_ref * mask / (1 - p)
            np.testing.assert_array_equal(mask_out.cpu().numpy(), mask)
        np.testing.assert_allclose(out.cpu().numpy(), o_ref, atol=2e-5)
        dy, db, dg, dbe = torch.empty_like(ty), torch.empty(H, device="cuda"), torch.empty(H, device="cuda"), \
            torch.empty(H, device="cuda")
        _kernels.check(lib.ppo_relu_ln_bwd(tdo.data_ptr(), h.data_ptr(), stats.data_ptr(), tg.data_ptr(),
                                           tm.data_ptr() if p > 0 else None, p, dy.data_ptr(), db.data_ptr(),
                                           dg.data_ptr(), dbe.data_ptr(), rows, H, _stream(torch)))
        d_in = dout.astype(np.float64) * (mask / (1 - p) if p > 0 else 1.0)
        dx_ref, _, db_ref, dg_ref, dbe_ref = po._block_backward(d_in, cache, np.eye(H), gamma.astype(np.float64))
        np.testing.assert_allclose(dy.cpu().numpy(), dx_ref, atol=3e-5)
        np.testing.assert_allclose(db.cpu().numpy(), db_ref, atol=2e-4)
        np.testing.assert_allclose(dg.cpu().numpy(), dg_ref, atol=2e-4)
        np.testing.assert_allclose(dbe.cpu().numpy(), dbe_ref, atol=2e-4)


@pytest.mark.parametrize("M,N,K", [(128, 256, 329), (128, 45, 256), (44, 1, 128), (300, 128, 256), (7, 256, 57)])
def test_small_m_gemms(torch_cuda, M, N, K):
    """The three GEMM kernels against float64 matmul on ragged shapes (nt: forward, nn: dA, tn: dW)."""
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels

    lib = _kernels.load()
    gcpu = torch.Generator().manual_seed(M * 1000 + N)
    a = torch.randn(M, K, generator=gcpu).cuda()
    w = torch.randn(N, K, generator=gcpu).cuda()
    dy = torch.randn(M, N, generator=gcpu).cuda()
    y, da, dw = torch.full((M, N), 7.0, device="cuda"), torch.full((M, K), 7.0, device="cuda"), \
        torch.full((N, K), 7.0, device="cuda")
    _kernels.check(lib.ppo_gemm_nt(a.data_ptr(), w.data_ptr(), y.data_ptr(), M, N, K, _stream(torch)))
    _kernels.check(lib.ppo_gemm_nn(dy.data_ptr(), w.data_ptr(), da.data_ptr(), M, N, K, _stream(torch)))
    _kernels.check(lib.ppo_gemm_tn(dy.data_ptr(), a.data_ptr(), dw.data_ptr(), M, N, K, _stream(torch)))
    a64, w64, dy64 = a.double(), w.double(), dy.double()
    for got, ref in ((y, a64 @ w64.t()), (da, dy64 @ w64), (dw, dy64.t() @ a64)):
        scale = ref.abs().max().item()
        assert (got.double() - ref).abs().max().item() < 2e-6 * scale * max(1.0, (max(M, N, K) / 64) ** 0.5)


def test_policy_head_matches_oracle(torch_cuda):
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels
    from oracle import ppo_oracle as po

    lib = _kernels.load()
    rs = np.random.RandomState(1)
    B, A = 128, 45
    z = rs.randn(B, A).astype(np.float32)
    b4 = (0.1 * rs.randn(A)).astype(np.float32)
    log_std = (-1 + 0.3 * rs.randn(A)).astype(np.float32)
    log_std[3] = 2.5  # outside the clamp: sigma = e^2, no gradient
    actions = (np.tanh(z + b4) + 0.4 * rs.randn(B, A)).astype(np.float32)
    adv = rs.randn(B).astype(np.float32)
    zz = (z + b4).astype(np.float64)
    _, _, logp0, _, _ = po.policy_head(zz, actions.astype(np.float64), np.zeros(B), adv.astype(np.float64),
                                       log_std.astype(np.float64), 0.2, 0.01)
    old = (logp0 + 0.15 * rs.randn(B)).astype(np.float32)  # ratios on both sides of the clip range
    loss, ent, logp, dz_ref, dls_ref = po.policy_head(zz, actions.astype(np.float64), old.astype(np.float64),
                                                      adv.astype(np.float64), log_std.astype(np.float64), 0.2, 0.01)
    d = lambda a: torch.as_tensor(a, device="cuda")  # noqa: E731
    ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device="cuda")
    dz, db4, dls, lp = torch.empty(B, A, device="cuda"), torch.empty(A, device="cuda"), torch.empty(A, device="cuda"), \
        torch.empty(B, device="cuda")
    tz, tb, tl, ta, to, tadv = d(z), d(b4), d(log_std), d(actions), d(old), d(adv)
    scratch = torch.zeros(130 * ((B + 7) // 8), device="cuda")
    ticket = torch.zeros(1, dtype=torch.int32, device="cuda")
    for _ in range(2):  # twice: the ticket must be left at zero
        _kernels.check(lib.ppo_policy_head(tz.data_ptr(), tb.data_ptr(), tl.data_ptr(), ta.data_ptr(), to.data_ptr(),
                                           tadv.data_ptr(), 0.2, 0.01, dz.data_ptr(), db4.data_ptr(), dls.data_ptr(),
                                           lp.data_ptr(), ctrl.data_ptr(), scratch.data_ptr(), ticket.data_ptr(), B, A,
                                           _stream(torch)))
    assert int(ticket.item()) == 0
    c = _kernels.PPOCtrl.from_buffer_copy(ctrl.cpu().numpy().tobytes())
    np.testing.assert_allclose(lp.cpu().numpy(), logp, rtol=2e-5, atol=2e-4)
    # a ratio within float32 rounding of the clip edge may fall on the other side: compare all but those rows
    ratio = np.exp(logp - old)
    safe = (np.abs(ratio - 0.8) > 1e-3) & (np.abs(ratio - 1.2) > 1e-3)
    np.testing.assert_allclose(dz.cpu().numpy()[safe], dz_ref[safe], atol=2e-5 * np.abs(dz_ref).max() + 1e-7)
    if safe.all():
        np.testing.assert_allclose(db4.cpu().numpy(), dz_ref.sum(0), atol=1e-4 * np.abs(dz_ref.sum(0)).max())
        np.testing.assert_allclose(dls.cpu().numpy(), dls_ref, atol=1e-4 * np.abs(dls_ref).max())
        assert abs(c.actor_loss - loss) < 1e-4 * max(1.0, abs(loss))
    assert dls.cpu().numpy()[3] == 0.0
    assert abs(c.entropy - ent) < 1e-5


@pytest.mark.parametrize("n", [2, 300, 4096, 70001])
def test_prepare_batch_matches_oracle(torch_cuda, n):
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels
    from oracle import ppo_oracle as po

    lib = _kernels.load()
    rs = np.random.RandomState(n)
    ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device="cuda")
    _kernels.check(lib.ppo_ctrl_init(ctrl.data_ptr(), 5, 1e-4, 2e-4, _stream(torch)))
    rms = po.RunningMeanStd()
    for it in range(2):  # twice: the running statistics carry over
        rewards = (2.0 + rs.randn(n)).astype(np.float32)
        dones = (rs.rand(n) < 0.05).astype(np.float32)
        values, next_values = rs.randn(n).astype(np.float32), rs.randn(n).astype(np.float32)
        d = lambda a: torch.as_tensor(a, device="cuda")  # noqa: E731
        tr, td, tv, tnv = d(rewards), d(dones), d(values), d(next_values)
        rn, ret, araw, adv = (torch.empty(n, device="cuda") for _ in range(4))
        _kernels.check(lib.ppo_prepare_batch(tr.data_ptr(), td.data_ptr(), tv.data_ptr(), tnv.data_ptr(), 0.99, 0.95,
                                             rn.data_ptr(), ret.data_ptr(), araw.data_ptr(), adv.data_ptr(),
                                             ctrl.data_ptr(), n, _stream(torch)))
        rn_ref = rms(rewards.astype(np.float64)).astype(np.float32).astype(np.float64)
        ret_ref, araw_ref, adv_ref = po.prepare_batch(rn_ref, dones.astype(np.float64), values.astype(np.float64),
                                                      next_values.astype(np.float64), 0.99)
        c = _kernels.PPOCtrl.from_buffer_copy(ctrl.cpu().numpy().tobytes())
        np.testing.assert_allclose([c.rms_mean, c.rms_var, c.rms_count], [rms.mean, rms.var, rms.count], rtol=1e-10)
        np.testing.assert_allclose(rn.cpu().numpy(), rn_ref, atol=1e-6)
        np.testing.assert_allclose(ret.cpu().numpy(), ret_ref, atol=1e-5)
        scale = np.abs(araw_ref).max()
        np.testing.assert_allclose(araw.cpu().numpy(), araw_ref, atol=2e-6 * scale + 1e-6)
        np.testing.assert_allclose(adv.cpu().numpy(), adv_ref, atol=2e-5)
        assert abs(c.mean_reward_n - rn_ref.mean()) < 1e-5


def test_sampling_and_permutation(torch_cuda):
    torch = torch_cuda
    from oracle import ppo_oracle as po

    ag = _agent(torch, 40, 45, seed=3)
    from diffusion_piano_b200.ppo import _kernels

    rs = np.random.RandomState(2)
    n = 4096
    states = torch.as_tensor(rs.randn(n, 40).astype(np.float32), device="cuda")
    ora = po.PPOOracle({k: v.numpy() for k, v in ag.actor_state_dict().items()},
                       {k: v.numpy() for k, v in ag.critic_state_dict().items()})
    mu, sigma = ora.act_mean_std(states.cpu().numpy())
    # explicit noise: exact formula
    noise = rs.randn(n, 45).astype(np.float32)
    a, lp = ag.select_actions(states, noise=torch.as_tensor(noise, device="cuda"))
    np.testing.assert_allclose(a.cpu().numpy(), mu + sigma * noise, atol=2e-5)
    np.testing.assert_allclose(lp.cpu().numpy(), ora.log_prob(states.cpu().numpy(), a.cpu().numpy().astype(np.float64)),
                               rtol=1e-4, atol=2e-3)
    # Philox noise: standard normal, fresh on every call, log-prob consistent with the drawn action
    a1, lp1 = ag.select_actions(states)
    a1, lp1 = a1.cpu().numpy().copy(), lp1.cpu().numpy().copy()
    a2, _ = ag.select_actions(states)
    eps = (a1 - mu) / sigma
    assert abs(eps.mean()) < 0.01 and abs(eps.std() - 1.0) < 0.01
    assert abs(((eps ** 3).mean())) < 0.05 and abs((eps ** 4).mean() - 3.0) < 0.1
    assert np.abs(a2.cpu().numpy() - a1).max() > 0.1
    np.testing.assert_allclose(lp1, ora.log_prob(states.cpu().numpy(), a1.astype(np.float64)), rtol=1e-4, atol=2e-3)
    # the minibatch permutation is a bijection of [0, N) and changes with the epoch
    lib = _kernels.load()
    perms = []
    for epoch in (0, 1):
        c = ag.read_ctrl()
        c.epoch = epoch
        tmp = torch.as_tensor(np.frombuffer(bytes(c), dtype=np.uint8).copy(), device="cuda")
        for nn in (300, 4096, 5000):
            out = torch.empty(nn, dtype=torch.int32, device="cuda")
            _kernels.check(lib.ppo_permutation(tmp.data_ptr(), 0, nn, out.data_ptr(), _stream(torch)))
            o = out.cpu().numpy()
            assert sorted(o.tolist()) == list(range(nn))
            if nn == 4096:
                perms.append(o)
                assert np.mean(o == np.arange(nn)) < 0.01
                # adjacent outputs are not adjacent inputs
                assert np.mean(np.abs(np.diff(o)) == 1) < 0.01
    assert np.mean(perms[0] == perms[1]) < 0.01


def _golden():
    from make_ppo_golden import ppo_fixture_inputs

    z = np.load(ROOT / "tests" / "golden" / "ppo_update.npz")
    state_dim, action_dim, n, batch, epochs, seed, stride = (int(v) for v in z["meta"])
    masks, i = [], 0
    while f"mask_{i}" in z:
        shp = tuple(z[f"mask_{i}_shape"])
        masks.append(np.unpackbits(z[f"mask_{i}"])[: shp[0] * shp[1]].reshape(shp))
        i += 1
    actor, critic, b = ppo_fixture_inputs(seed, state_dim, action_dim, n)
    return z, (state_dim, action_dim, n, batch, epochs, stride), masks, actor, critic, b


def test_update_matches_reference_and_oracle(torch_cuda):
    """The full update on the reference's recorded permutations and dropout masks: against the reference's own
    parameters after the update (golden) and against the oracle's intermediate quantities."""
    torch = torch_cuda
    from oracle import ppo_oracle as po

    z, (sd, ad, n, batch, epochs, stride), masks, actor, critic, b = _golden()
    ag = _agent(torch, sd, ad, actor, critic, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch, ppo_epochs=epochs)
    n_mb = epochs * ((n + batch - 1) // batch)
    dm = lambda m: torch.as_tensor(m.astype(np.uint8), device="cuda")  # noqa: E731
    bb = ag.update(b["states"], b["actions"], b["rewards"].astype(np.float32), z["log_probs"], b["next_states"],
                   b["dones"], perms=z["perms"], value_masks=[dm(m) for m in masks[0:2]],
                   next_value_masks=[dm(m) for m in masks[2:4]],
                   minibatch_masks=[[dm(m) for m in masks[4 + 2 * i: 6 + 2 * i]] for i in range(n_mb)])
    ora = po.PPOOracle(actor, critic, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch, ppo_epochs=epochs)
    f = [m.astype(np.float64) for m in masks]
    inter = ora.update(b["states"], b["actions"], b["rewards"].astype(np.float32), z["log_probs"], b["next_states"],
                       b["dones"], perms=z["perms"], value_masks=f[0:2], next_value_masks=f[2:4],
                       minibatch_masks=[f[4 + 2 * i: 6 + 2 * i] for i in range(n_mb)])
    np.testing.assert_allclose(bb["packed"][2].cpu().numpy(), inter["values"], atol=2e-5)
    np.testing.assert_allclose(bb["packed"][3].cpu().numpy(), inter["next_values"], atol=2e-5)
    np.testing.assert_allclose(bb["returns"].cpu().numpy(), inter["returns"], atol=2e-5)
    np.testing.assert_allclose(bb["adv"].cpu().numpy(), inter["advantages"], atol=1e-4)
    for name, before, mine, theirs in (("actor", actor, ag.actor_state_dict(), ora.actor_state()),
                                       ("critic", critic, ag.critic_state_dict(), ora.critic_state())):
        for k, v in mine.items():
            v = v.numpy().astype(np.float64)
            b0 = before[k].astype(np.float64)
            # against the oracle, every element
            np.testing.assert_allclose(v - b0, theirs[k] - b0, atol=4e-6, err_msg=f"oracle {name}/{k}")
            # against the reference's own result
            ref = z[f"{name}/{k}/sample"].astype(np.float64)
            np.testing.assert_allclose((v - b0).ravel()[::stride], ref - b0.ravel()[::stride], atol=4e-6,
                                       err_msg=f"reference {name}/{k}")
    c = ag.read_ctrl()
    np.testing.assert_allclose([c.rms_mean, c.rms_var, c.rms_count], z["rms"], rtol=1e-6)
    assert c.adam_t[0] == n_mb and c.adam_t[1] == n_mb
    al, cl, ent = ora.log[-1]
    assert abs(c.critic_loss - cl) < 1e-4 * max(1.0, cl) and abs(c.entropy - ent) < 1e-5
    assert abs(c.actor_loss - al) < 2e-4 * max(1.0, abs(al))


def test_fused_value_head_matches_unfused(torch_cuda):
    """ppo_value_head_fused (the critic's last Linear and its dA inside the head kernel) against the three-launch
    path: same update on the reference's recorded permutations and masks."""
    torch = torch_cuda
    z, (sd, ad, n, batch, epochs, stride), masks, actor, critic, b = _golden()
    n_mb = epochs * ((n + batch - 1) // batch)
    dm = lambda m: torch.as_tensor(m.astype(np.uint8), device="cuda")  # noqa: E731
    results = []
    for fused in (False, True):
        ag = _agent(torch, sd, ad, actor, critic, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch, ppo_epochs=epochs,
                    fuse_value_head=fused)
        ag.update(b["states"], b["actions"], b["rewards"].astype(np.float32), z["log_probs"], b["next_states"],
                  b["dones"], perms=z["perms"], value_masks=[dm(m) for m in masks[0:2]],
                  next_value_masks=[dm(m) for m in masks[2:4]],
                  minibatch_masks=[[dm(m) for m in masks[4 + 2 * i: 6 + 2 * i]] for i in range(n_mb)])
        results.append((ag.actor.theta.clone(), ag.critic.theta.clone(), ag.read_ctrl()))
    for i in (0, 1):
        assert (results[0][i] - results[1][i]).abs().max().item() < 4e-6
    assert abs(results[0][2].actor_loss - results[1][2].actor_loss) < 1e-5
    assert abs(results[0][2].critic_loss - results[1][2].critic_loss) < 1e-5


def test_graph_update_matches_oracle(torch_cuda):
    """The CUDA-graph path (device-side permutation, no dropout so that the oracle can follow): parameters after
    two updates against the oracle fed the materialised permutations."""
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels
    from make_ppo_golden import ppo_fixture_inputs
    from oracle import ppo_oracle as po

    lib = _kernels.load()
    sd, ad, n, batch, epochs = 57, 45, 600, 128, 3
    actor, critic, b = ppo_fixture_inputs(11, sd, ad, n)
    ag = _agent(torch, sd, ad, actor, critic, batch_size=batch, ppo_epochs=epochs, dropout_p=0.0, seed=5)
    ora = po.PPOOracle(actor, critic, batch_size=batch, ppo_epochs=epochs)
    logp = ora.log_prob(b["states"], b["actions"].astype(np.float64)).astype(np.float32)
    rs = np.random.RandomState(3)
    for it in range(2):
        lp = (logp + 0.1 * rs.randn(n)).astype(np.float32)
        rewards = (1.0 + rs.randn(n)).astype(np.float32)
        perms = []
        c = ag.read_ctrl()
        for e in range(epochs):
            c.epoch = it * epochs + e
            tmp = torch.as_tensor(np.frombuffer(bytes(c), dtype=np.uint8).copy(), device="cuda")
            out = torch.empty(n, dtype=torch.int32, device="cuda")
            _kernels.check(lib.ppo_permutation(tmp.data_ptr(), 0, n, out.data_ptr(), _stream(torch)))
            perms.append(out.cpu().numpy())
        launches0 = ag.kernel_launches
        ag.update(b["states"], b["actions"], rewards, lp, b["next_states"], b["dones"])
        assert ag.kernel_launches - launches0 > epochs * 5 * 20  # the kernels inside the replayed graph are counted
        ora.update(b["states"], b["actions"], rewards, lp, b["next_states"], b["dones"], perms=perms)
        torch.cuda.synchronize()
        assert ag.read_ctrl().epoch == (it + 1) * epochs
    for name, before, mine, theirs in (("actor", actor, ag.actor_state_dict(), ora.actor_state()),
                                       ("critic", critic, ag.critic_state_dict(), ora.critic_state())):
        for k, v in mine.items():
            b0 = before[k].astype(np.float64)
            # 30 Adam steps of 1e-4 / 2e-4 (displacement up to 6e-3): m / sqrt(v) turns float32 rounding of a tiny
            # gradient into a fraction of a full step, so a few elements drift by multiples of the learning rate
            # while the bulk agrees to a few 1e-6
            err = np.abs(v.numpy().astype(np.float64) - theirs[k]).ravel()
            assert np.median(err) < 3e-6 and np.quantile(err, 0.98) < 5e-5 and err.max() < 2e-3, \
                (name, k, np.median(err), np.quantile(err, 0.98), err.max())


def test_dropout_statistics_and_determinism(torch_cuda):
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels

    lib = _kernels.load()
    ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device="cuda")
    _kernels.check(lib.ppo_ctrl_init(ctrl.data_ptr(), 9, 1e-4, 2e-4, _stream(torch)))
    rows, H = 4096, 256
    y = torch.randn(rows, H, device="cuda")
    ones, zeros = torch.ones(H, device="cuda"), torch.zeros(H, device="cuda")
    out, m1, m2, m3 = torch.empty_like(y), *(torch.empty(rows, H, dtype=torch.uint8, device="cuda") for _ in range(3))

    def run(mask, stream_id):
        _kernels.check(lib.ppo_relu_ln_fwd(y.data_ptr(), zeros.data_ptr(), ones.data_ptr(), zeros.data_ptr(),
                                           out.data_ptr(), None, None, rows, H, 0.1, None, mask.data_ptr(),
                                           ctrl.data_ptr(), stream_id, _stream(torch)))

    run(m1, 1)
    run(m2, 1)
    assert torch.equal(m1, m2)  # same (seed, counter, stream): same mask
    keep = m1.float().mean().item()
    assert abs(keep - 0.9) < 2e-3
    assert (out[m1 == 0] == 0).all()
    run(m3, 2)
    assert 0.75 < (m1 == m3).float().mean().item() < 0.9  # independent streams agree with probability 0.82
    _kernels.check(lib.ppo_advance(ctrl.data_ptr(), 0, 0, _stream(torch)))
    run(m2, 1)
    assert 0.75 < (m1 == m2).float().mean().item() < 0.9  # new counter: new mask


def test_plateau_kernel_matches_torch(torch_cuda):
    torch = torch_cuda
    from diffusion_piano_b200.ppo import _kernels
    from oracle import ppo_oracle as po

    lib = _kernels.load()
    ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device="cuda")
    _kernels.check(lib.ppo_ctrl_init(ctrl.data_ptr(), 0, 1e-4, 2e-4, _stream(torch)))
    sa, sc = po.ReduceLROnPlateau("max", 0.5, 3), po.ReduceLROnPlateau("min", 0.5, 3)
    lra, lrc = 1e-4, 2e-4
    rs = np.random.RandomState(4)
    for i in range(40):
        c = _kernels.PPOCtrl.from_buffer_copy(ctrl.cpu().numpy().tobytes())
        c.mean_reward_n = float(np.float32(-1.0 / (1 + i // 8) + 0.01 * rs.rand()))
        c.critic_loss = float(np.float32(1.0 / (1 + i // 12) + 0.01 * rs.rand()))
        ctrl.copy_(torch.as_tensor(np.frombuffer(bytes(c), dtype=np.uint8).copy()))
        _kernels.check(lib.ppo_plateau_step(ctrl.data_ptr(), 0.5, 3, _stream(torch)))
        lra, lrc = sa.step(c.mean_reward_n, lra), sc.step(c.critic_loss, lrc)
        got = _kernels.PPOCtrl.from_buffer_copy(ctrl.cpu().numpy().tobytes())
        np.testing.assert_allclose([got.lr[0], got.lr[1]], [lra, lrc], rtol=1e-6)
    assert lra < 1e-4 and lrc < 2e-4  # both schedulers fired


def test_train_loop_config4_smoke(torch_cuda):
    """Config 4 in small: 'Crossing Field Cut 10s' (no usable fingering -> OT reward, obs without the fingering
    block), the reference's loop for a few control steps; rewards stay finite, parameters move, and the
    checkpoint round-trips."""
    torch = torch_cuda
    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.ppo import TrainConfig, train

    task = suite.PianoWithShadowHands(music.load("CrossingFieldCut10s"), n_steps_lookahead=1, trim_silence=True)
    env = suite.BatchedEnvironment(task, 256, canonical_actions=True, auto_reset=True)
    cfg = TrainConfig(num_envs=256, num_episodes=1, batch_size=128, ppo_epochs=2, max_steps_per_episode=4)
    agent, hist = train(env, cfg, log=lambda s: None)
    st = agent.stats()
    assert st["adam_steps"] == 4 * 2 * 2
    assert np.isfinite(hist["mean_rewards"][0]) and np.isfinite(st["actor_loss"]) and np.isfinite(st["critic_loss"])
    import tempfile

    with tempfile.TemporaryDirectory() as d:
        agent.checkpoint_dir = Path(d)
        path = agent.save_checkpoint(0, hist)
        from diffusion_piano_b200.ppo import PPOAgent

        other = PPOAgent(env.obs_dim, env.action_dim, device="cuda:0", seed=99, checkpoint_dir=d)
        other.load_checkpoint(path)
        assert torch.equal(other.actor.theta, agent.actor.theta) and torch.equal(other.critic.m, agent.critic.m)
        # optimiser, scheduler, normaliser and RNG position are restored; the seed stays the loading rank's own
        mine, theirs = other.read_ctrl(), agent.read_ctrl()
        for f in ("lr", "adam_t", "best", "num_bad"):
            assert list(getattr(mine, f)) == list(getattr(theirs, f)), f
        for f in ("rms_mean", "rms_var", "rms_count", "rng_counter", "epoch"):
            assert getattr(mine, f) == getattr(theirs, f), f
        assert mine.seed != theirs.seed
    env.close()


def test_reference_loop_with_numpy_api(torch_cuda):
    """The reference's loop body as written (parallelized_base_v2.py:116-166): numpy observation dicts from
    VectorizedPianoEnv, hand-concatenated states, numpy in / numpy out agent calls."""
    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.ppo import PPOAgent

    num_envs = 8
    vec_env = suite.VectorizedPianoEnv(num_envs, music.load("GurenNoYumiyaCut14s"))
    observation_spec = vec_env.envs[0].observation_spec()
    action_spec = vec_env.envs[0].action_spec()
    state_dim = sum(int(np.prod(spec.shape)) for spec in observation_spec.values())
    action_dim = action_spec.shape[0]
    assert (state_dim, action_dim) == (329, 45)
    agent = PPOAgent(state_dim=state_dim, action_dim=action_dim, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=4,
                     device="cuda", checkpoint_dir="/tmp/ppo_numpy_api", ppo_epochs=2)
    keys = ("goal", "fingering", "piano/state", "piano/sustain_state", "rh_shadow_hand/joints_pos",
            "lh_shadow_hand/joints_pos")

    def flat(o):
        return np.stack([np.concatenate([o[k][i].flatten() for k in keys]) for i in range(num_envs)])

    obs = vec_env.reset()
    episode_rewards = np.zeros(num_envs)
    before = agent.actor.theta.clone()
    for _ in range(3):
        states = flat(obs)
        actions, log_probs = agent.select_actions(states)
        assert isinstance(actions, np.ndarray) and actions.shape == (num_envs, 45) and log_probs.shape == (num_envs,)
        next_obs, rewards, dones = vec_env.step(actions)
        next_states = flat(next_obs)
        episode_rewards += rewards
        agent.update(states=states, actions=actions, rewards=rewards, log_probs=log_probs, next_states=next_states,
                     dones=dones)
        assert not all(dones)
        obs = next_obs
    st = agent.stats()
    assert st["adam_steps"] == 3 * 2 * 2 and np.isfinite(st["actor_loss"]) and np.isfinite(episode_rewards).all()
    assert (agent.actor.theta - before).abs().max().item() > 0
    vec_env.batched.close()


def test_large_minibatch_takes_the_cublas_path(torch_cuda):
    """batch_size above the small-M limit (512): cuBLAS GEMMs and the unfused heads; same update rule."""
    torch = torch_cuda
    from make_ppo_golden import ppo_fixture_inputs
    from oracle import ppo_oracle as po

    sd, ad, n, batch = 33, 45, 1500, 1000
    actor, critic, b = ppo_fixture_inputs(21, sd, ad, n)
    ag = _agent(torch, sd, ad, actor, critic, batch_size=batch, ppo_epochs=1, dropout_p=0.0, seed=2, use_graphs=False)
    assert batch > ag.small_m
    ora = po.PPOOracle(actor, critic, batch_size=batch, ppo_epochs=1)
    logp = (ora.log_prob(b["states"], b["actions"].astype(np.float64)) + 0.05).astype(np.float32)
    perm = np.random.RandomState(1).permutation(n)[None].astype(np.int32)
    rewards = b["rewards"].astype(np.float32)
    ag.update(b["states"], b["actions"], rewards, logp, b["next_states"], b["dones"], perms=perm)
    ora.update(b["states"], b["actions"], rewards, logp, b["next_states"], b["dones"], perms=perm)
    for mine, theirs in ((ag.actor_state_dict(), ora.actor_state()), (ag.critic_state_dict(), ora.critic_state())):
        for k, v in mine.items():
            np.testing.assert_allclose(v.numpy(), theirs[k], atol=5e-6, err_msg=k)


def test_checkpoint_is_the_reference_layout_and_round_trips(tmp_path):
    """`save_checkpoint` writes what the reference's `ppo_v2.PPOAgent.load_checkpoint`
    (ppo_v2.py:327-336) reads - same top-level keys, state_dict keys and order, torch Adam / ReduceLROnPlateau
    state_dicts (layout recorded from the reference itself: tests/golden/ppo_checkpoint_layout.json,
    tools/make_ppo_checkpoint_layout.py) - torch's own load_state_dict accepts it, and loading it back restores
    moments, step counts, learning rates and scheduler state while the rank-specific seed stays the loader's."""
    import json
    import torch
    from diffusion_piano_b200.ppo import PPOAgent
    lay = json.loads((ROOT / "tests" / "golden" / "ppo_checkpoint_layout.json").read_text())
    S, A, N = lay["state_dim"], lay["action_dim"], 64
    sys.path.insert(0, str(ROOT / "tools"))
    from make_ppo_golden import ppo_fixture_inputs
    _, _, b = ppo_fixture_inputs(3, S, A, N)
    agent = PPOAgent(S, A, batch_size=32, ppo_epochs=1, device="cuda:0", seed=5, checkpoint_dir=tmp_path, use_graphs=False)
    for _ in range(2):
        agent.update(b["states"], b["actions"], b["rewards"].astype(np.float32), b["log_probs"], b["next_states"], b["dones"])
    path = agent.save_checkpoint(7, {"mean_rewards": [1.0]})
    ck = torch.load(path, map_location="cpu", weights_only=False)
    assert list(ck)[:len(lay["keys"])] == lay["keys"] and set(ck) - set(lay["keys"]) == {"b200_state"}
    for net in ("actor", "critic"):
        assert [[k, list(v.shape)] for k, v in ck[f"{net}_state_dict"].items()] == lay[f"{net}_state_dict"]
        osd, ol = ck[f"{net}_optimizer_state_dict"], lay[f"{net}_optimizer"]
        g = osd["param_groups"][0]
        assert len(osd["param_groups"]) == ol["n_groups"] and sorted(g) == ol["group_keys"] and g["params"] == ol["params"]
        assert list(g["betas"]) == ol["betas"] and g["eps"] == ol["eps"] and g["weight_decay"] == ol["weight_decay"]
        assert [list(osd["state"][i]["exp_avg"].shape) for i in g["params"]] == ol["shapes"]
        st = osd["state"][g["params"][0]]
        assert sorted(st) == ol["state_keys"] and torch.is_tensor(st["step"]) and float(st["step"]) == 4.0  # 2 updates x 2 minibatches
        assert sorted(ck[f"{net}_scheduler_state_dict"]) == lay[f"{net}_scheduler_keys"]
        for k in ("mode", "factor", "patience"):
            assert ck[f"{net}_scheduler_state_dict"][k] == lay[f"{net}_scheduler"][k]
    # what the reference's load_checkpoint does with the file: torch modules of its architecture + load_state_dict
    h = 256
    seq = lambda dims, drop: torch.nn.Sequential(*[m for i in range(3) for m in (
        [torch.nn.Linear(dims[i], dims[i + 1]), torch.nn.ReLU(), torch.nn.LayerNorm(dims[i + 1])]
        + ([torch.nn.Dropout(0.1)] if drop[i] else []))] + [torch.nn.Linear(dims[3], dims[4])])

    class Actor(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.network = torch.nn.Sequential(*list(seq((S, h, h, h, A), (0, 0, 0))), torch.nn.Tanh())
            self.log_std = torch.nn.Parameter(torch.ones(A) * -1.0)

    class Critic(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.network = seq((S, h, h, h // 2, 1), (1, 1, 0))

    ref_a, ref_c = Actor(), Critic()
    ref_a.load_state_dict(ck["actor_state_dict"]); ref_c.load_state_dict(ck["critic_state_dict"])
    for mod, net, mode in ((ref_a, "actor", "max"), (ref_c, "critic", "min")):
        opt = torch.optim.Adam(mod.parameters(), lr=1e-4, betas=(0.9, 0.999), eps=1e-5)
        sch = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode=mode, factor=0.5, patience=100)
        opt.load_state_dict(ck[f"{net}_optimizer_state_dict"]); sch.load_state_dict(ck[f"{net}_scheduler_state_dict"])
        mine = agent.actor if net == "actor" else agent.critic
        for name, prm in mod.named_parameters():
            np.testing.assert_array_equal(opt.state[prm]["exp_avg"].numpy(), mine._view(mine.m, name).cpu().numpy())
            np.testing.assert_array_equal(opt.state[prm]["exp_avg_sq"].numpy(), mine._view(mine.v, name).cpu().numpy())
            np.testing.assert_array_equal(prm.detach().numpy(), mine.p[name].cpu().numpy())
    # and back: a fresh agent of another "rank" resumes with identical optimiser / scheduler state, its own seed
    c0 = agent.read_ctrl()
    other = PPOAgent(S, A, batch_size=32, ppo_epochs=1, device="cuda:0", seed=9, checkpoint_dir=tmp_path, use_graphs=False)
    seed_before = other.read_ctrl().seed
    episode, rewards = other.load_checkpoint(path)
    c1 = other.read_ctrl()
    assert episode == 7 and rewards == {"mean_rewards": [1.0]}
    assert c1.seed == seed_before != c0.seed
    for f in ("lr", "adam_t", "best", "num_bad"):
        assert list(getattr(c1, f)) == list(getattr(c0, f)), f
    assert (c1.rms_mean, c1.rms_var, c1.rms_count, c1.rng_counter) == (c0.rms_mean, c0.rms_var, c0.rms_count, c0.rng_counter)
    for mine, theirs in ((agent.actor, other.actor), (agent.critic, other.critic)):
        for buf in ("theta", "m", "v"):
            assert torch.equal(getattr(mine, buf), getattr(theirs, buf)), buf
    # the next update from the restored state gives the same parameters as continuing the original (same seed needed:
    # compare the deterministic part only - Adam step counts and learning rates advance alike)
    other.update(b["states"], b["actions"], b["rewards"].astype(np.float32), b["log_probs"], b["next_states"], b["dones"])
    assert list(other.read_ctrl().adam_t) == [6, 6]
