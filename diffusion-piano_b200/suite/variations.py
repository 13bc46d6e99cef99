"""MIDI augmentations applied at the start of an episode (reference `robopianist/suite/variations.py:27-184`).

Same classes, constructor arguments and call signature (`var(initial_value=midi, random_state=rs)`) as the
reference; dm_control's `variation.Variation` / `distributions` are absent here, so the draws are made directly on
the `numpy.random.RandomState` in the order the reference makes them (the probability draw first, then the value):
`UniformChoice` -> `random_state.choice`, `Uniform(lo, hi)` -> `random_state.uniform(lo, hi)`.
A task built with `augmentations=[...]` keeps K augmented songs resident on the GPU (a "song bank") and every env
takes one of them at its reset: `suite/environment.py`, `pp_set_song_bank` / `pp_set_song_ids`.
"""

from __future__ import annotations

from typing import Sequence

import numpy as np

from .. import music
from ..music import constants, midi_file


class Variation:
    """Stand-in for `dm_control.composer.variation.base.Variation`: a callable returning the new value."""

    def __call__(self, initial_value=None, current_value=None, random_state=None):  # pragma: no cover
        raise NotImplementedError


def _require_midi(initial_value):
    if initial_value is None or not isinstance(initial_value, midi_file.MidiFile):
        raise ValueError("Expected `initial_value` to be provided and be a midi_file.MidiFile.")
    return initial_value


class MidiSelect(Variation):
    """Randomly select a MIDI file from the registry (variations.py:27-46)."""

    def __init__(self, midi_names: Sequence[str] = ()) -> None:
        self._midi_names = list(midi_names)

    def __call__(self, initial_value=None, current_value=None, random_state=None) -> midi_file.MidiFile:
        del initial_value, current_value
        random = random_state or np.random
        return music.load(str(random.choice(self._midi_names)))


class MidiTemporalStretch(Variation):
    """With probability `prob`, stretch by a factor drawn from U(1 - range, 1 + range) (variations.py:49-83)."""

    def __init__(self, prob: float, stretch_range: float) -> None:
        self._prob = prob
        self._range = stretch_range

    def __call__(self, initial_value=None, current_value=None, random_state=None) -> midi_file.MidiFile:
        del current_value
        random = random_state or np.random
        if random.uniform(0.0, 1.0) > self._prob:
            return _require_midi(initial_value)
        stretch_factor = 1.0 + random.uniform(-self._range, self._range)
        return initial_value.stretch(stretch_factor)


def _pitch_bounds(midi: midi_file.MidiFile, limit: int):
    pitches = [note.pitch for note in midi.seq.notes]
    low = max(constants.MIN_MIDI_PITCH_PIANO - min(pitches), -limit)
    high = min(constants.MAX_MIDI_PITCH_PIANO - max(pitches), limit)
    return low, high


class MidiPitchShift(Variation):
    """With probability `prob`, transpose by an integer number of semitones in [-range, range], truncated so that
    no note leaves the piano (variations.py:86-134)."""

    def __init__(self, prob: float, shift_range: int) -> None:
        self._prob = prob
        if not isinstance(shift_range, int):
            raise ValueError("`shift_range` must be an integer.")
        self._shift_range = shift_range

    def __call__(self, initial_value=None, current_value=None, random_state=None) -> midi_file.MidiFile:
        del current_value
        random = random_state or np.random
        if random.uniform(0.0, 1.0) > self._prob:
            return _require_midi(initial_value)
        if self._shift_range == 0:
            return initial_value
        low, high = _pitch_bounds(initial_value, self._shift_range)
        shift = random.randint(low, high + 1)
        if shift == 0:
            return initial_value
        return initial_value.transpose(int(shift))


class MidiOctaveShift(Variation):
    """With probability `prob`, transpose by whole octaves in [-range, range], truncated to the piano
    (variations.py:137-184)."""

    def __init__(self, prob: float, octave_range: int) -> None:
        self._prob = prob
        if not isinstance(octave_range, int):
            raise ValueError("`octave_range` must be an integer.")
        self._octave_range = octave_range

    def __call__(self, initial_value=None, current_value=None, random_state=None) -> midi_file.MidiFile:
        del current_value
        random = random_state or np.random
        if random.uniform(0.0, 1.0) > self._prob:
            return _require_midi(initial_value)
        if self._octave_range == 0:
            return initial_value
        low, high = _pitch_bounds(initial_value, self._octave_range * 12)
        shift = random.randint(low // 12, high // 12 + 1)
        if shift == 0:
            return initial_value
        return initial_value.transpose(int(shift) * 12)
