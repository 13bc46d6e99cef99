"""Musical metrics (reference robopianist/wrappers/evaluation.py:30-177).

The reference wrapper records `task.piano.activation` after every step and, at the end of an episode,
runs sklearn's `precision_recall_fscore_support(average="binary", zero_division=1)` per step against the
song's notes, then averages over the steps. Here the step kernel accumulates the same per-step numbers
on the device (`pp_get_metrics`); this wrapper only keeps the reference's deque-of-episodes interface.
"""

from __future__ import annotations

from collections import deque
from typing import Deque, Dict, NamedTuple, Sequence

import numpy as np


class EpisodeMetrics(NamedTuple):
    precision: float
    recall: float
    f1: float


_NAMES = ("precision", "recall", "f1", "sustain_precision", "sustain_recall", "sustain_f1")


class MidiEvaluationWrapper:
    """Wraps a single dm_env-style `suite.Environment`; same methods as the reference wrapper."""

    def __init__(self, environment, deque_size: int = 1) -> None:
        self._environment = environment
        self._episodes: Deque[np.ndarray] = deque(maxlen=deque_size)

    def __getattr__(self, name):
        return getattr(self._environment, name)

    def step(self, action):
        timestep = self._environment.step(action)
        if timestep.last():
            self._episodes.append(self._environment.batched.episode_metrics()[0][0])
        return timestep

    def reset(self):
        return self._environment.reset()

    def get_musical_metrics(self) -> Dict[str, float]:
        """Mean precision / recall / F1 over the last `deque_size` episodes."""
        if not self._episodes:
            raise ValueError("No episode metrics available yet.")

        def _mean(seq: Sequence[float]) -> float:
            return sum(seq) / len(seq)

        return {name: _mean([float(e[j]) for e in self._episodes]) for j, name in enumerate(_NAMES)}

    def get_last_episode(self) -> EpisodeMetrics:
        e = self._episodes[-1]
        return EpisodeMetrics(float(e[0]), float(e[1]), float(e[2]))
