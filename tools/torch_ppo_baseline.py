"""Plain-PyTorch restatement of the reference's PPO v2 update (ppo_v2.py:51-105, 211-309) used ONLY as the timed
comparison arm of tools/bench_ppo.py (`--impl torch`): nn.Sequential networks, autograd, torch.optim.Adam,
clip_grad_norm_, a Python loop over minibatches with one `.item()` per minibatch as the reference's wandb
logging forces, and the Python loop over the batch index for the "GAE". /root/reference is not on the GPU box
and its module needs wandb, so it cannot be imported there; this file follows it statement by statement."""
import numpy as np
import torch
import torch.nn as nn


class Actor(nn.Module):
    def __init__(self, s, a, h=256):
        super().__init__()
        self.network = nn.Sequential(nn.Linear(s, h), nn.ReLU(), nn.LayerNorm(h), nn.Linear(h, h), nn.ReLU(),
                                     nn.LayerNorm(h), nn.Linear(h, h), nn.ReLU(), nn.LayerNorm(h), nn.Linear(h, a),
                                     nn.Tanh())
        self.log_std = nn.Parameter(torch.ones(a) * -1.0)

    def forward(self, x):
        return torch.distributions.Normal(self.network(x), torch.exp(torch.clamp(self.log_std, -20, 2)))


class Critic(nn.Module):
    def __init__(self, s, h=256):
        super().__init__()
        self.network = nn.Sequential(nn.Linear(s, h), nn.ReLU(), nn.LayerNorm(h), nn.Dropout(0.1), nn.Linear(h, h),
                                     nn.ReLU(), nn.LayerNorm(h), nn.Dropout(0.1), nn.Linear(h, h // 2), nn.ReLU(),
                                     nn.LayerNorm(h // 2), nn.Linear(h // 2, 1))
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.orthogonal_(m.weight, gain=0.01)
                nn.init.constant_(m.bias, 0)

    def forward(self, x):
        return self.network(x)


class TorchPPO:
    def __init__(self, s, a, lr=1e-4, gamma=0.99, epsilon=0.2, entropy_coef=0.01, max_grad_norm=0.5, ppo_epochs=10,
                 batch_size=128, device="cuda", python_gae=True):
        self.device = torch.device(device)
        self.actor, self.critic = Actor(s, a).to(self.device), Critic(s).to(self.device)
        self.aopt = torch.optim.Adam(self.actor.parameters(), lr=lr, eps=1e-5)
        self.copt = torch.optim.Adam(self.critic.parameters(), lr=lr * 2, eps=1e-5)
        self.gamma, self.epsilon, self.entropy_coef, self.max_grad_norm = gamma, epsilon, entropy_coef, max_grad_norm
        self.ppo_epochs, self.batch_size, self.python_gae = ppo_epochs, batch_size, python_gae
        self.mean, self.var, self.count = 0.0, 1.0, 1e-4

    def select_actions(self, states):
        with torch.no_grad():
            dist = self.actor(states)
            actions = dist.sample()
            return actions, dist.log_prob(actions).sum(1)

    def update(self, states, actions, rewards, log_probs, next_states, dones):
        r = rewards.double().cpu().numpy()  # the reference normalises in numpy on the host
        bm, bv, bc = r.mean(), r.var(), r.shape[0]
        delta, tot = bm - self.mean, self.count + bc
        self.var = (self.var * self.count + bv * bc + delta * delta * self.count * bc / tot) / tot
        self.mean, self.count = self.mean + delta * bc / tot, tot
        rewards = torch.as_tensor((r - self.mean) / np.sqrt(self.var + 1e-8), dtype=torch.float32, device=self.device)
        dones = dones.float()
        with torch.no_grad():
            values = self.critic(states).squeeze(-1)
            next_values = self.critic(next_states).squeeze(-1)
            returns = rewards + self.gamma * next_values * (1 - dones)
            advantages = torch.zeros_like(rewards)
            if self.python_gae:  # as written in the reference: a Python loop over the batch index
                gae = 0
                n = len(rewards)
                for t in reversed(range(n)):
                    nv = next_values[t] if t == n - 1 else values[t + 1]
                    d = rewards[t] + self.gamma * nv * (1 - dones[t]) - values[t]
                    gae = d + self.gamma * 0.95 * (1 - dones[t]) * gae
                    advantages[t] = gae
            else:  # same recursion on the host in numpy (a favourable variant of the baseline)
                v, nv_, rr, dd = (x.cpu().numpy() for x in (values, next_values, rewards, dones))
                adv = np.zeros_like(rr)
                gae = 0.0
                for t in range(len(rr) - 1, -1, -1):
                    nv = nv_[t] if t == len(rr) - 1 else v[t + 1]
                    gae = rr[t] + self.gamma * nv * (1 - dd[t]) - v[t] + self.gamma * 0.95 * (1 - dd[t]) * gae
                    adv[t] = gae
                advantages = torch.as_tensor(adv, device=self.device)
        advantages = (advantages - advantages.mean()) / (advantages.std() + 1e-8)
        n = states.shape[0]
        logged = 0.0
        for _ in range(self.ppo_epochs):
            perm = torch.randperm(n, device=self.device)
            for s in range(0, n, self.batch_size):
                idx = perm[s:s + self.batch_size]
                dist = self.actor(states[idx])
                new_lp = dist.log_prob(actions[idx]).sum(1)
                entropy = dist.entropy().mean()
                ratio = torch.exp(new_lp - log_probs[idx])
                surr1 = ratio * advantages[idx]
                surr2 = torch.clamp(ratio, 1 - self.epsilon, 1 + self.epsilon) * advantages[idx]
                actor_loss = -torch.min(surr1, surr2).mean() - self.entropy_coef * entropy
                critic_loss = nn.functional.mse_loss(self.critic(states[idx]).squeeze(-1), returns[idx])
                self.copt.zero_grad()
                critic_loss.backward()
                torch.nn.utils.clip_grad_norm_(self.critic.parameters(), self.max_grad_norm)
                self.copt.step()
                self.aopt.zero_grad()
                actor_loss.backward()
                torch.nn.utils.clip_grad_norm_(self.actor.parameters(), self.max_grad_norm)
                self.aopt.step()
                logged += actor_loss.item()  # the reference's wandb.log reads six scalars per minibatch
        return logged
