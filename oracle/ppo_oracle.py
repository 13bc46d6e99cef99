"""CPU oracle of the reference's PPO v2 agent (TEST INFRASTRUCTURE ONLY - never imported by the product).

numpy float64 restatement, with hand-derived backward passes, of `/root/reference/ppo_v2.py`:
  Actor            ppo_v2.py:51-74    (Linear -> ReLU -> LayerNorm) x3 -> Linear -> Tanh, Normal(mean, exp(clamp(log_std)))
  Critic           ppo_v2.py:76-105   (Linear -> ReLU -> LayerNorm [-> Dropout 0.1]) x3 -> Linear
  RunningMeanStd   ppo_v2.py:107-131
  PPOAgent.update  ppo_v2.py:211-309  reward normalisation, returns, the index-wise "GAE" recursion, advantage
                                      normalisation, shuffled minibatches, critic step then actor step,
                                      clip_grad_norm_ + Adam(eps=1e-5), ReduceLROnPlateau bookkeeping
  select_actions   ppo_v2.py:202-209

Randomness (the DataLoader permutations and the critic's dropout masks) is an INPUT here, so that the
oracle can be pinned against a run of the reference itself with recorded randomness
(`tools/make_ppo_golden.py` -> `tests/golden/ppo_update.npz`, checked by `tests/test_ppo_oracle.py`) and the
CUDA update can be checked against the oracle on the same permutations and masks.

Parameter naming follows the reference's state_dict: `network.0.weight` ... and `log_std`.
"""

from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import numpy as np

LN_EPS = 1e-5  # torch.nn.LayerNorm default
ACTOR_LINEARS = (0, 3, 6, 9)
ACTOR_NORMS = (2, 5, 8)
CRITIC_LINEARS = (0, 4, 8, 11)
CRITIC_NORMS = (2, 6, 10)
CRITIC_DROPOUT_AFTER = (True, True, False)  # ppo_v2.py:83,88 - none after the third block
DROPOUT_P = 0.1


def _block_forward(x, W, b, g, be):
    """Linear -> ReLU -> LayerNorm. Returns output and the cache the backward needs."""
    h = np.maximum(x @ W.T + b, 0.0)
    mu = h.mean(axis=1, keepdims=True)
    var = ((h - mu) ** 2).mean(axis=1, keepdims=True)  # biased, as LayerNorm
    rstd = 1.0 / np.sqrt(var + LN_EPS)
    xhat = (h - mu) * rstd
    return xhat * g + be, (x, h, xhat, rstd)


def _block_backward(dout, cache, W, g):
    x, h, xhat, rstd = cache
    dg = (dout * xhat).sum(axis=0)
    dbe = dout.sum(axis=0)
    dxhat = dout * g
    dh = rstd * (dxhat - dxhat.mean(axis=1, keepdims=True) - xhat * (dxhat * xhat).mean(axis=1, keepdims=True))
    dy = np.where(h > 0.0, dh, 0.0)
    return dy @ W, dy.T @ x, dy.sum(axis=0), dg, dbe


class MLP:
    """Three Linear/ReLU/LayerNorm blocks and a final Linear, parameters keyed as the reference's state_dict."""

    def __init__(self, params: Dict[str, np.ndarray], linears: Sequence[int], norms: Sequence[int],
                 dropout_after: Sequence[bool] = (False, False, False)):
        self.p = {k: np.asarray(v, np.float64).copy() for k, v in params.items()}
        self.linears, self.norms, self.dropout_after = linears, norms, dropout_after

    def forward(self, x, masks: Optional[List[np.ndarray]] = None):
        """masks: one keep-mask [B, H] (0/1) per dropout layer, None = eval (no dropout)."""
        self.caches, self.masks = [], masks
        a, mi = np.asarray(x, np.float64), 0
        for blk in range(3):
            li, ni = self.linears[blk], self.norms[blk]
            a, cache = _block_forward(a, self.p[f"network.{li}.weight"], self.p[f"network.{li}.bias"],
                                      self.p[f"network.{ni}.weight"], self.p[f"network.{ni}.bias"])
            self.caches.append(cache)
            if self.dropout_after[blk] and masks is not None:
                a = a * masks[mi] / (1.0 - DROPOUT_P)
                mi += 1
        self.last_in = a
        li = self.linears[3]
        return a @ self.p[f"network.{li}.weight"].T + self.p[f"network.{li}.bias"]

    def backward(self, dz) -> Dict[str, np.ndarray]:
        grads = {}
        li = self.linears[3]
        grads[f"network.{li}.weight"] = dz.T @ self.last_in
        grads[f"network.{li}.bias"] = dz.sum(axis=0)
        da = dz @ self.p[f"network.{li}.weight"]
        mi = sum(self.dropout_after) - 1 if self.masks is not None else -1
        for blk in (2, 1, 0):
            li, ni = self.linears[blk], self.norms[blk]
            if self.dropout_after[blk] and self.masks is not None:
                da = da * self.masks[mi] / (1.0 - DROPOUT_P)
                mi -= 1
            da, dW, db, dg, dbe = _block_backward(da, self.caches[blk], self.p[f"network.{li}.weight"],
                                                  self.p[f"network.{ni}.weight"])
            grads[f"network.{li}.weight"], grads[f"network.{li}.bias"] = dW, db
            grads[f"network.{ni}.weight"], grads[f"network.{ni}.bias"] = dg, dbe
        return grads


class Adam:
    """torch.optim.Adam(betas=(0.9, 0.999), eps=1e-5), no weight decay / amsgrad (ppo_v2.py:157-168)."""

    def __init__(self, params: Dict[str, np.ndarray], lr: float, eps: float = 1e-5):
        self.lr, self.eps, self.b1, self.b2, self.t = lr, eps, 0.9, 0.999, 0
        self.m = {k: np.zeros_like(v) for k, v in params.items()}
        self.v = {k: np.zeros_like(v) for k, v in params.items()}

    def step(self, params, grads):
        self.t += 1
        bc1, bc2 = 1.0 - self.b1 ** self.t, 1.0 - self.b2 ** self.t
        for k, g in grads.items():
            self.m[k] = self.b1 * self.m[k] + (1 - self.b1) * g
            self.v[k] = self.b2 * self.v[k] + (1 - self.b2) * g * g
            params[k] -= (self.lr / bc1) * self.m[k] / (np.sqrt(self.v[k]) / np.sqrt(bc2) + self.eps)


def clip_grad_norm(grads: Dict[str, np.ndarray], max_norm: float) -> float:
    """torch.nn.utils.clip_grad_norm_: coef = max_norm / (total + 1e-6), clamped to 1."""
    total = float(np.sqrt(sum(float((g * g).sum()) for g in grads.values())))
    coef = min(max_norm / (total + 1e-6), 1.0)
    for k in grads:
        grads[k] = grads[k] * coef
    return total


class ReduceLROnPlateau:
    """torch.optim.lr_scheduler.ReduceLROnPlateau(mode, factor=0.5, patience=100), threshold 1e-4 'rel',
    cooldown 0, min_lr 0, eps 1e-8 (ppo_v2.py:171-184)."""

    def __init__(self, mode: str, factor: float = 0.5, patience: int = 100):
        self.mode, self.factor, self.patience, self.threshold = mode, factor, patience, 1e-4
        self.best = np.inf if mode == "min" else -np.inf
        self.num_bad = 0

    def step(self, metric: float, lr: float) -> float:
        if self.mode == "min":
            better = metric < self.best * (1.0 - self.threshold)
        else:
            better = metric > self.best * (1.0 + self.threshold)
        if better:
            self.best, self.num_bad = metric, 0
        else:
            self.num_bad += 1
        if self.num_bad > self.patience:
            new_lr = lr * self.factor
            self.num_bad = 0
            if lr - new_lr > 1e-8:
                return new_lr
        return lr


class RunningMeanStd:
    """ppo_v2.py:107-131."""

    def __init__(self, epsilon: float = 1e-4):
        self.mean, self.var, self.count = 0.0, 1.0, epsilon

    def __call__(self, x):
        x = np.asarray(x, np.float64)
        bm, bv, bc = x.mean(), x.var(), x.shape[0]
        delta = bm - self.mean
        tot = self.count + bc
        new_mean = self.mean + delta * bc / tot
        m2 = self.var * self.count + bv * bc + delta * delta * self.count * bc / tot
        self.mean, self.var, self.count = new_mean, m2 / tot, tot
        return (x - self.mean) / np.sqrt(self.var + 1e-8)


def prepare_batch(rewards_n, dones, values, next_values, gamma: float, lam: float = 0.95):
    """ppo_v2.py:224-249: returns, the recursion over the BATCH index (the reference runs its "GAE" over the
    env index of one control step, not over time - restated as written), advantage normalisation with the
    unbiased std."""
    n = len(rewards_n)
    returns = rewards_n + gamma * next_values * (1.0 - dones)
    adv = np.zeros(n)
    gae = 0.0
    for t in range(n - 1, -1, -1):
        nv = next_values[t] if t == n - 1 else values[t + 1]
        delta = rewards_n[t] + gamma * nv * (1.0 - dones[t]) - values[t]
        gae = delta + gamma * lam * (1.0 - dones[t]) * gae
        adv[t] = gae
    adv_n = (adv - adv.mean()) / (adv.std(ddof=1) + 1e-8)
    return returns, adv, adv_n


def policy_head(z, actions, old_logp, adv, log_std, epsilon: float, entropy_coef: float):
    """Tanh mean, Normal log-prob / entropy, clipped surrogate (ppo_v2.py:262-271) and its gradient with
    respect to the pre-tanh output z and log_std. torch.min splits the gradient evenly on ties, and inside the
    clip range surr1 == surr2, so the derivative there is the unclipped one."""
    B, A = z.shape
    mu = np.tanh(z)
    ls = np.clip(log_std, -20.0, 2.0)
    sigma = np.exp(ls)
    diff = actions - mu
    logp = (-(diff ** 2) / (2 * sigma ** 2) - ls - 0.5 * np.log(2 * np.pi)).sum(axis=1)
    entropy = (0.5 + 0.5 * np.log(2 * np.pi) + ls).mean() * np.ones(())  # mean over [B, A] of a row-constant
    ratio = np.exp(logp - old_logp)
    surr1 = ratio * adv
    clipped = np.clip(ratio, 1 - epsilon, 1 + epsilon)
    surr2 = clipped * adv
    loss = -np.minimum(surr1, surr2).mean() - entropy_coef * entropy
    inside = (ratio >= 1 - epsilon) & (ratio <= 1 + epsilon)
    w1 = np.where(surr1 < surr2, 1.0, np.where(surr1 == surr2, 0.5, 0.0))
    w2 = np.where(surr2 < surr1, 1.0, np.where(surr1 == surr2, 0.5, 0.0))
    dsurr = w1 * adv * ratio + w2 * adv * ratio * inside  # d/dlogp of min(surr1, surr2)
    dlogp = -dsurr / B
    dmu = dlogp[:, None] * diff / sigma ** 2
    dz = dmu * (1.0 - mu ** 2)
    live = (log_std >= -20.0) & (log_std <= 2.0)
    dls = (dlogp[:, None] * (diff ** 2 / sigma ** 2 - 1.0)).sum(axis=0) - entropy_coef / A
    return float(loss), float(entropy), logp, dz, np.where(live, dls, 0.0)


class PPOOracle:
    def __init__(self, actor_params, critic_params, lr=1e-4, gamma=0.99, epsilon=0.2, entropy_coef=0.01,
                 max_grad_norm=0.5, ppo_epochs=10, batch_size=64):
        self.actor = MLP({k: v for k, v in actor_params.items() if k != "log_std"}, ACTOR_LINEARS, ACTOR_NORMS)
        self.log_std = np.asarray(actor_params["log_std"], np.float64).copy()
        self.critic = MLP(critic_params, CRITIC_LINEARS, CRITIC_NORMS, CRITIC_DROPOUT_AFTER)
        self.actor_params = dict(self.actor.p)
        self.actor_params["log_std"] = self.log_std
        self.actor_opt = Adam(self.actor_params, lr)
        self.critic_opt = Adam(self.critic.p, lr * 2)
        self.actor_sched = ReduceLROnPlateau("max")
        self.critic_sched = ReduceLROnPlateau("min")
        self.normalizer = RunningMeanStd()
        self.gamma, self.epsilon, self.entropy_coef = gamma, epsilon, entropy_coef
        self.max_grad_norm, self.ppo_epochs, self.batch_size = max_grad_norm, ppo_epochs, batch_size
        self.log = []

    def act_mean_std(self, states):
        return np.tanh(self.actor.forward(states)), np.exp(np.clip(self.log_std, -20.0, 2.0))

    def log_prob(self, states, actions):
        mu, sigma = self.act_mean_std(states)
        return (-((actions - mu) ** 2) / (2 * sigma ** 2) - np.log(sigma) - 0.5 * np.log(2 * np.pi)).sum(axis=1)

    def update(self, states, actions, rewards, log_probs, next_states, dones, perms, value_masks=None,
               next_value_masks=None, minibatch_masks=None):
        """perms: [epochs][N] index permutations. *_masks: dropout keep-masks (two per critic forward), None =
        dropout off. Returns the intermediate quantities for the tests."""
        states, next_states = np.asarray(states, np.float64), np.asarray(next_states, np.float64)
        actions, old_logp = np.asarray(actions, np.float64), np.asarray(log_probs, np.float64)
        dones = np.asarray(dones, np.float64)
        # the reference casts the normalised rewards to float32 (torch.FloatTensor, ppo_v2.py:213)
        rewards_n = self.normalizer(rewards).astype(np.float32).astype(np.float64)
        values = self.critic.forward(states, value_masks)[:, 0]
        next_values = self.critic.forward(next_states, next_value_masks)[:, 0]
        returns, adv_raw, adv = prepare_batch(rewards_n, dones, values, next_values, self.gamma)
        n, bs = len(rewards_n), self.batch_size
        mb = 0
        critic_loss = 0.0
        for e in range(self.ppo_epochs):
            perm = np.asarray(perms[e])
            for s in range(0, n, bs):
                idx = perm[s:s + bs]
                B = len(idx)
                z = self.actor.forward(states[idx])
                aloss, ent, _, dz, dls = policy_head(z, actions[idx], old_logp[idx], adv[idx], self.log_std,
                                                     self.epsilon, self.entropy_coef)
                masks = None if minibatch_masks is None else minibatch_masks[mb]
                v = self.critic.forward(states[idx], masks)[:, 0]
                critic_loss = float(((v - returns[idx]) ** 2).mean())
                cg = self.critic.backward((2.0 * (v - returns[idx]) / B)[:, None])
                clip_grad_norm(cg, self.max_grad_norm)
                self.critic_opt.step(self.critic.p, cg)
                ag = self.actor.backward(dz)
                ag["log_std"] = dls
                clip_grad_norm(ag, self.max_grad_norm)
                # the Adam step must see the same arrays the networks use
                self.actor_opt.step(self.actor_params, ag)
                self.log.append((aloss, critic_loss, ent))
                mb += 1
        self.actor_opt.lr = self.actor_sched.step(float(rewards_n.mean()), self.actor_opt.lr)
        self.critic_opt.lr = self.critic_sched.step(critic_loss, self.critic_opt.lr)
        return dict(rewards_n=rewards_n, values=values, next_values=next_values, returns=returns,
                    advantages_raw=adv_raw, advantages=adv)

    def actor_state(self):
        return {k: v.copy() for k, v in self.actor_params.items()}

    def critic_state(self):
        return {k: v.copy() for k, v in self.critic.p.items()}
