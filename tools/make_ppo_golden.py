"""Generates tests/golden/ppo_update.npz by running the REFERENCE's own `ppo_v2.PPOAgent` (imported from
/root/reference in this container, CPU) on seeded inputs, with its randomness recorded.

What is stubbed, and why it does not change the algorithm:
* `wandb` is not installed: an empty module object stands in (the agent is built with use_wandb=False).
* `torch.randperm` (DataLoader shuffle) and `torch.nn.functional.dropout` (critic) are wrapped so that the
  permutations and keep-masks they draw are RECORDED; the dropout wrapper draws Bernoulli(1-p) and scales by
  1/(1-p) exactly as torch's does.
* `ReduceLROnPlateau(verbose=True)`: the keyword no longer exists in this torch; a shim drops it.
Parameters and inputs come from `ppo_fixture_inputs(seed)` below (numpy RandomState), so the fixture only
stores the recorded randomness and the reference's results (a strided sample of every parameter tensor after
the update plus per-tensor sums).

Run here (not on the GPU box): python tools/make_ppo_golden.py
"""
import sys
import types
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
STATE_DIM, ACTION_DIM, N, BATCH, EPOCHS, SEED, STRIDE = 40, 45, 300, 128, 2, 7, 16


def _shapes(state_dim, action_dim, hidden=256):
    actor = {"log_std": (action_dim,)}
    for li, (i, o) in zip((0, 3, 6, 9), ((state_dim, hidden), (hidden, hidden), (hidden, hidden), (hidden, action_dim))):
        actor[f"network.{li}.weight"], actor[f"network.{li}.bias"] = (o, i), (o,)
    for ni in (2, 5, 8):
        actor[f"network.{ni}.weight"], actor[f"network.{ni}.bias"] = (hidden,), (hidden,)
    critic = {}
    dims = ((state_dim, hidden), (hidden, hidden), (hidden, hidden // 2), (hidden // 2, 1))
    for li, (i, o) in zip((0, 4, 8, 11), dims):
        critic[f"network.{li}.weight"], critic[f"network.{li}.bias"] = (o, i), (o,)
    for ni, h in zip((2, 6, 10), (hidden, hidden, hidden // 2)):
        critic[f"network.{ni}.weight"], critic[f"network.{ni}.bias"] = (h,), (h,)
    return actor, critic


def ppo_fixture_inputs(seed=SEED, state_dim=STATE_DIM, action_dim=ACTION_DIM, n=N):
    """Deterministic parameters and one batch of transitions (float32), shared by the generator and the tests."""
    rs = np.random.RandomState(seed)

    def fill(shapes):
        out = {}
        for k in sorted(shapes):
            shp = shapes[k]
            if k == "log_std":
                v = -1.0 + 0.1 * rs.randn(*shp)
            elif len(shp) == 2:
                v = rs.randn(*shp) / np.sqrt(shp[1])
            elif k.endswith("weight"):
                v = 1.0 + 0.1 * rs.randn(*shp)
            else:
                v = 0.1 * rs.randn(*shp)
            out[k] = v.astype(np.float32)
        return out

    a_shapes, c_shapes = _shapes(state_dim, action_dim)
    actor, critic = fill(a_shapes), fill(c_shapes)
    batch = dict(
        states=rs.randn(n, state_dim).astype(np.float32),
        next_states=rs.randn(n, state_dim).astype(np.float32),
        actions=(0.5 * rs.randn(n, action_dim)).astype(np.float32),
        rewards=(1.0 + rs.randn(n)).astype(np.float64),
        log_probs=(-20.0 + 3.0 * rs.randn(n)).astype(np.float32),
        dones=(rs.rand(n) < 0.1),
    )
    return actor, critic, batch


def main():
    import torch

    sys.modules.setdefault("wandb", types.ModuleType("wandb"))
    sys.path.insert(0, "/root/reference")
    real_plateau = torch.optim.lr_scheduler.ReduceLROnPlateau

    class Plateau(real_plateau):
        def __init__(self, *a, verbose=None, **k):
            super().__init__(*a, **k)

    torch.optim.lr_scheduler.ReduceLROnPlateau = Plateau
    import ppo_v2  # the reference, unmodified

    actor, critic, batch = ppo_fixture_inputs()
    # old log-probs must be consistent with the policy for a meaningful ratio: take them from the reference actor
    agent = ppo_v2.PPOAgent(STATE_DIM, ACTION_DIM, lr=1e-4, gamma=0.99, epsilon=0.2, batch_size=BATCH,
                            ppo_epochs=EPOCHS, device="cpu", checkpoint_dir="/tmp/ppo_golden_ckpt", use_wandb=False)
    agent.actor.load_state_dict({k: torch.from_numpy(v) for k, v in actor.items()})
    agent.critic.load_state_dict({k: torch.from_numpy(v) for k, v in critic.items()})
    with torch.no_grad():
        dist = agent.actor(torch.from_numpy(batch["states"]))
        ref_logp = dist.log_prob(torch.from_numpy(batch["actions"])).sum(1).numpy()
    # perturb so that some ratios leave the clip range on both sides
    rs = np.random.RandomState(SEED + 1)
    log_probs = (ref_logp + 0.15 * rs.randn(N)).astype(np.float32)

    perms, masks = [], []
    real_randperm, real_dropout = torch.randperm, torch.nn.functional.dropout

    def rec_randperm(n, *a, **k):
        p = real_randperm(n, *a, **k)
        perms.append(p.numpy().copy())
        return p

    def rec_dropout(input, p=0.5, training=True, inplace=False):
        if not training or p == 0.0:
            return input
        keep = torch.empty_like(input).bernoulli_(1.0 - p)
        masks.append(keep.numpy().astype(bool))
        return input * keep / (1.0 - p)

    torch.manual_seed(SEED)
    torch.randperm, torch.nn.functional.dropout = rec_randperm, rec_dropout
    try:
        agent.update(states=batch["states"], actions=batch["actions"], rewards=batch["rewards"],
                     log_probs=log_probs, next_states=batch["next_states"], dones=batch["dones"])
    finally:
        torch.randperm, torch.nn.functional.dropout = real_randperm, real_dropout

    # RandomSampler draws a second permutation per epoch and keeps none of it (num_samples % n == 0)
    perms = perms[::2]
    n_mb = EPOCHS * ((N + BATCH - 1) // BATCH)
    assert len(perms) == EPOCHS, len(perms)
    assert len(masks) == 4 + 2 * n_mb, len(masks)
    out = {"perms": np.stack(perms).astype(np.int32), "log_probs": log_probs,
           "meta": np.array([STATE_DIM, ACTION_DIM, N, BATCH, EPOCHS, SEED, STRIDE])}
    for i, m in enumerate(masks):
        out[f"mask_{i}_shape"] = np.array(m.shape)
        out[f"mask_{i}"] = np.packbits(m.ravel())
    for name, net in (("actor", agent.actor), ("critic", agent.critic)):
        for k, v in net.state_dict().items():
            a = v.numpy().astype(np.float64).ravel()
            out[f"{name}/{k}/sample"] = a[::STRIDE].astype(np.float32)
            out[f"{name}/{k}/sum"] = np.array([a.sum(), np.abs(a).sum()])
        # .grad still holds the clipped gradient of the LAST minibatch: pins every backward formula
        for k, v in net.named_parameters():
            g = v.grad.numpy().astype(np.float64).ravel()
            out[f"{name}/{k}/grad_sample"] = g[::STRIDE].astype(np.float32)
            out[f"{name}/{k}/grad_sum"] = np.array([g.sum(), np.abs(g).sum()])
    out["rms"] = np.array([agent.reward_normalizer.mean, agent.reward_normalizer.var, agent.reward_normalizer.count])
    out["lrs"] = np.array([agent.actor_optimizer.param_groups[0]["lr"], agent.critic_optimizer.param_groups[0]["lr"]])
    dst = ROOT / "tests" / "golden" / "ppo_update.npz"
    np.savez_compressed(dst, **out)
    print("wrote", dst, dst.stat().st_size, "bytes;", len(masks), "masks,", len(perms), "perms")


if __name__ == "__main__":
    main()
