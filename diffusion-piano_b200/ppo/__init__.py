"""Device-resident PPO v2 (reference ppo_v2.py + the training loop of parallelized_base_v2.py:68-188)."""
from .agent import PPOAgent
from .train import TrainConfig, train

__all__ = ["PPOAgent", "TrainConfig", "train"]
