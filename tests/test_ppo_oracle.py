"""The PPO oracle (oracle/ppo_oracle.py) pinned against a recorded run of the reference's own ppo_v2.PPOAgent.update
(tests/golden/ppo_update.npz, made by tools/make_ppo_golden.py from /root/reference/ppo_v2.py:211-309)."""
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))

from make_ppo_golden import ppo_fixture_inputs  # noqa: E402
from oracle import ppo_oracle  # noqa: E402


def load_golden():
    z = np.load(ROOT / "tests" / "golden" / "ppo_update.npz")
    meta = z["meta"]
    masks = []
    i = 0
    while f"mask_{i}" in z:
        shp = tuple(z[f"mask_{i}_shape"])
        masks.append(np.unpackbits(z[f"mask_{i}"])[: shp[0] * shp[1]].reshape(shp).astype(np.float64))
        i += 1
    return z, meta, masks


def run_oracle_on_golden():
    z, meta, masks = load_golden()
    state_dim, action_dim, n, batch, epochs, seed, stride = (int(v) for v in meta)
    actor, critic, b = ppo_fixture_inputs(seed, state_dim, action_dim, n)
    ora = ppo_oracle.PPOOracle(actor, critic, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch, ppo_epochs=epochs)
    n_mb = epochs * ((n + batch - 1) // batch)
    mb_masks = [masks[4 + 2 * i: 6 + 2 * i] for i in range(n_mb)]
    ora.last_grads = {}
    inter = ora.update(b["states"], b["actions"], b["rewards"], z["log_probs"], b["next_states"], b["dones"],
                       perms=z["perms"], value_masks=masks[0:2], next_value_masks=masks[2:4],
                       minibatch_masks=mb_masks)
    return z, stride, actor, critic, ora, inter


def test_oracle_matches_reference_update():
    z, stride, actor0, critic0, ora, _ = run_oracle_on_golden()
    for name, before, after in (("actor", actor0, ora.actor_state()), ("critic", critic0, ora.critic_state())):
        for k, v in after.items():
            ref = z[f"{name}/{k}/sample"].astype(np.float64)
            d_ref = ref - before[k].astype(np.float64).ravel()[::stride]
            d_ora = (v - before[k].astype(np.float64)).ravel()[::stride]
            # 6 Adam steps of 1e-4 / 2e-4: the parameter displacement agrees to a few % of one step; the
            # reference's float32 parameters resolve ~1e-8 of it
            np.testing.assert_allclose(d_ora, d_ref, atol=4e-6, err_msg=f"{name}/{k}")
            assert abs(v.sum() - z[f"{name}/{k}/sum"][0]) < 1e-4 * max(1.0, z[f"{name}/{k}/sum"][1])
    np.testing.assert_allclose([ora.normalizer.mean, ora.normalizer.var, ora.normalizer.count], z["rms"], rtol=1e-12)
    np.testing.assert_allclose([ora.actor_opt.lr, ora.critic_opt.lr], z["lrs"])


def test_oracle_matches_reference_gradients():
    """The clipped gradients of the last minibatch (what .grad holds after update) against the oracle's."""
    z, meta, masks = load_golden()
    state_dim, action_dim, n, batch, epochs, seed, stride = (int(v) for v in meta)
    actor, critic, b = ppo_fixture_inputs(seed, state_dim, action_dim, n)
    ora = ppo_oracle.PPOOracle(actor, critic, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=batch, ppo_epochs=epochs)
    grads = {}
    real_clip = ppo_oracle.clip_grad_norm

    def spy(g, max_norm):
        total = real_clip(g, max_norm)
        grads["critic" if "network.11.weight" in g else "actor"] = {k: v.copy() for k, v in g.items()}
        return total

    ppo_oracle.clip_grad_norm = spy
    try:
        n_mb = epochs * ((n + batch - 1) // batch)
        ora.update(b["states"], b["actions"], b["rewards"], z["log_probs"], b["next_states"], b["dones"],
                   perms=z["perms"], value_masks=masks[0:2], next_value_masks=masks[2:4],
                   minibatch_masks=[masks[4 + 2 * i: 6 + 2 * i] for i in range(n_mb)])
    finally:
        ppo_oracle.clip_grad_norm = real_clip
    for name in ("actor", "critic"):
        for k, g in grads[name].items():
            ref = z[f"{name}/{k}/grad_sample"].astype(np.float64)
            scale = np.abs(ref).max() + 1e-12
            np.testing.assert_allclose(g.ravel()[::stride] / scale, ref / scale, atol=2e-3, err_msg=f"{name}/{k}")
            gs = z[f"{name}/{k}/grad_sum"]
            assert abs(g.sum() - gs[0]) <= 2e-3 * gs[1] + 1e-9, (name, k)


def test_prepare_batch_scan_identity():
    """The index-wise recursion of ppo_v2.py:236-246 is a linear recurrence: composing (a, b) pairs in any
    association gives the same result - the property the CUDA block scan relies on."""
    rs = np.random.RandomState(0)
    n = 1000
    r, v, nv = rs.randn(n), rs.randn(n), rs.randn(n)
    d = (rs.rand(n) < 0.05).astype(np.float64)
    _, adv, _ = ppo_oracle.prepare_batch(r, d, v, nv, 0.99)
    nxt = np.concatenate([v[1:], nv[-1:]])
    delta = r + 0.99 * nxt * (1 - d) - v
    c = 0.99 * 0.95 * (1 - d)
    # chunked evaluation: summaries (A, B) per chunk, then a carry from the right
    out = np.zeros(n)
    carry = 0.0
    for lo in range(n - 100, -1, -100):
        g = carry
        for t in range(lo + 99, lo - 1, -1):
            g = delta[t] + c[t] * g
            out[t] = g
        carry = g
    np.testing.assert_allclose(out, adv, rtol=1e-12)


def test_plateau_matches_torch():
    import torch

    opt = torch.optim.SGD([torch.nn.Parameter(torch.zeros(1))], lr=1.0)
    sch = torch.optim.lr_scheduler.ReduceLROnPlateau(opt, mode="min", factor=0.5, patience=3)
    mine = ppo_oracle.ReduceLROnPlateau("min", 0.5, 3)
    lr = 1.0
    rs = np.random.RandomState(1)
    for i in range(60):
        m = float(1.0 / (1 + i // 10) + 0.01 * rs.rand())
        sch.step(m)
        lr = mine.step(m, lr)
        assert lr == opt.param_groups[0]["lr"], i
