"""The reference's training loop (`parallelized_base_v2.py:68-188`) with every tensor resident in HBM.

Per control step, exactly as the reference: flatten observations -> `agent.select_actions` -> `vec_env.step`
-> `agent.update(states, actions, rewards, log_probs, next_states, dones)` (ppo_v2 updates on the batch of ONE
control step across the envs) -> stop the episode when all envs are done. Differences: the observation is
already the flat [N, D] row the reference concatenates by hand (same key order, :122-131), nothing is copied
to the host inside an episode, and the episode length is known (all envs play the same song in lockstep), so
the `all(dones)` test costs one device read per episode instead of one per step.
"""

from __future__ import annotations

import json
import time
from dataclasses import dataclass
from pathlib import Path
from typing import Callable, List, Optional

import torch

from . import dist
from .agent import PPOAgent


@dataclass
class TrainConfig:
    num_envs: int = 2          # parallelized_base_v2.py:70
    num_episodes: int = 20000  # :71
    lr: float = 1e-4           # :100-107
    gamma: float = 0.99
    epsilon: float = 0.2
    batch_size: int = 128
    ppo_epochs: int = 10       # PPOAgent default (ppo_v2.py:145)
    checkpoint_frequency: int = 1000
    checkpoint_dir: Optional[str] = None
    model_dir: Optional[str] = None
    seed: int = 0
    max_steps_per_episode: Optional[int] = None  # benches cut episodes short; None = until all envs are done


def train(env, cfg: TrainConfig, process_group=None, agent: Optional[PPOAgent] = None,
          on_step: Optional[Callable[[int, int], None]] = None, log: Callable[[str], None] = print):
    """env: a `suite.BatchedEnvironment` built with canonical_actions=True and auto_reset=True.
    Returns (agent, history) with history as the reference's dict of lists."""
    stamp = time.strftime("%Y%m%d_%H%M%S")
    if agent is None:
        agent = PPOAgent(state_dim=env.obs_dim, action_dim=env.action_dim, lr=cfg.lr, gamma=cfg.gamma,
                         epsilon=cfg.epsilon, batch_size=cfg.batch_size, ppo_epochs=cfg.ppo_epochs,
                         device=env.device, checkpoint_dir=cfg.checkpoint_dir or f"checkpoints/{stamp}",
                         seed=cfg.seed, process_group=process_group)
    rank = agent.rank
    history = {"episode_rewards": [], "mean_rewards": []}
    n = env.num_envs
    episode_rewards = torch.zeros(n, dtype=torch.float32, device=env.device)
    states = torch.empty((n, env.obs_dim), dtype=torch.float32, device=env.device)
    frames = env.task.song.n_frames
    for episode in range(cfg.num_episodes):
        obs = env.reset_device()
        episode_rewards.zero_()
        step = 0
        while True:
            states.copy_(obs)  # the env reuses its observation buffer
            actions, log_probs = agent.select_actions(states)
            next_obs, rewards, dones, _, _ = env.step_device(actions)
            episode_rewards += rewards
            agent.update(states=states, actions=actions, rewards=rewards, log_probs=log_probs, next_states=next_obs,
                         dones=dones)
            step += 1
            if on_step is not None:
                on_step(episode, step)
            if cfg.max_steps_per_episode is not None and step >= cfg.max_steps_per_episode:
                break
            # all envs play the same song in lockstep: they are done together after `frames` steps, so the
            # reference's `if all(dones): break` needs a device read only then (or every step when a wrong press
            # may end an env early)
            # With several ranks the decision is global (every rank evaluates `may_be_done` alike, so all of them
            # enter the all-reduce together): a rank leaving alone would strand the others in update's collectives.
            # `step % frames == 0` also catches envs that ended early, auto-reset and are done again later.
            # (with MIDI augmentations every env has its own episode length: look every step)
            may_be_done = (step % frames == 0 or bool(getattr(env.task, "_wrong_press_termination", False))
                           or bool(getattr(env, "_augment", False)))
            if may_be_done:
                done_here = bool(dones.all())
                if process_group is not None:
                    done_here = dist.all_ranks_agree(done_here, env.device, process_group)
                if done_here:
                    break
            obs = next_obs
        # statistics of the episode: one small all-reduce outside the step
        total = torch.stack([episode_rewards.sum(), torch.tensor(float(n), device=env.device)])
        if process_group is not None:
            torch.distributed.all_reduce(total, group=process_group)
        mean_reward = float(total[0] / total[1])
        if rank == 0:
            log(f"Episode {episode}, Mean Reward: {mean_reward}")
        history["episode_rewards"].append(episode_rewards.cpu().tolist())
        history["mean_rewards"].append(mean_reward)
        if rank == 0 and episode > 0 and episode % cfg.checkpoint_frequency == 0:
            agent.save_checkpoint(episode, history)
    if rank == 0 and cfg.model_dir is not None:
        model_dir = Path(cfg.model_dir)
        model_dir.mkdir(parents=True, exist_ok=True)
        agent.save_model(model_dir / f"final_model_{stamp}.pt")
        with open(model_dir / f"training_history_{stamp}.json", "w") as f:
            json.dump(history, f)
    return agent, history
