"""RoboPianist-style suite entry points (reference robopianist/suite/__init__.py:26-93)."""

from pathlib import Path
from typing import Any, Mapping, Optional, Union

from .. import music
from .environment import BatchedEnvironment, Environment, VectorizedPianoEnv
from .tasks import HandSide, PianoWithOneShadowHand, PianoWithShadowHands, RewardType, SelfActuatedPiano

_BASE_REPERTOIRE_NAME = "RoboPianist-repertoire-150-{}-v0"
_DEBUG_BASE_NAME = "RoboPianist-debug-{}-v0"
DEBUG = [_DEBUG_BASE_NAME.format(name) for name in music.DEBUG_MIDIS]
ALL = list(DEBUG)
_ENV_NAME_TO_MIDI = {_DEBUG_BASE_NAME.format(name): name for name in music.DEBUG_MIDIS}


def load(environment_name: str, midi_file: Optional[Union[str, Path]] = None, seed: Optional[int] = None,
         stretch: float = 1.0, shift: int = 0, recompile_physics: bool = False, legacy_step: bool = True,
         task_kwargs: Optional[Mapping[str, Any]] = None, device: int = 0) -> Environment:
    """Same contract as reference suite.load (:50-93); returns a dm_env-style Environment."""
    if midi_file is not None:
        midi = music.load(midi_file, stretch=stretch, shift=shift)
    else:
        if environment_name not in ALL:
            raise ValueError(f"Unknown environment {environment_name}. Available environments: {ALL}")
        midi = music.load(_ENV_NAME_TO_MIDI[environment_name], stretch=stretch, shift=shift)
    task = PianoWithShadowHands(midi=midi, **dict(task_kwargs or {}))
    return Environment(task, device=device, random_state=seed, recompile_physics=recompile_physics,
                       legacy_step=legacy_step)


__all__ = ["ALL", "DEBUG", "load", "BatchedEnvironment", "Environment", "VectorizedPianoEnv",
           "PianoWithShadowHands", "SelfActuatedPiano", "RewardType"]
