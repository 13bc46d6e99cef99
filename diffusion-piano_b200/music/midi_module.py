"""MIDI events from the piano's key activations (reference robopianist/models/piano/midi_module.py:27-140 and
robopianist/music/midi_message.py).

The reference's MidiModule runs in `Piano.after_substep` after every physics substep: a key that became active
emits NoteOn(note, velocity 127, time), one that became inactive NoteOff, the pedal SustainOn / SustainOff, all
stamped with `physics.data.time`. Here the step kernel records the activation bits after every substep
(`pp_set_substep_activations`) and this class replays them with the same rules, methods and callbacks."""

from __future__ import annotations

import enum
from dataclasses import dataclass
from typing import Callable, List, Optional, Union

import numpy as np

NUM_KEYS = 88
MIN_MIDI_PITCH_PIANO = 21  # key 0 = A0 (midi_file.key_number_to_midi_number)


@enum.unique
class EventType(enum.Enum):
    NOTE_ON = enum.auto()
    NOTE_OFF = enum.auto()
    SUSTAIN_ON = enum.auto()
    SUSTAIN_OFF = enum.auto()


@dataclass
class NoteOn:
    note: int
    velocity: int
    time: float
    event_type = EventType.NOTE_ON


@dataclass
class NoteOff:
    note: int
    time: float
    event_type = EventType.NOTE_OFF


@dataclass
class SustainOn:
    time: float
    event_type = EventType.SUSTAIN_ON


@dataclass
class SustainOff:
    time: float
    event_type = EventType.SUSTAIN_OFF


MidiMessage = Union[NoteOn, NoteOff, SustainOn, SustainOff]


def unpack_activation_words(words: np.ndarray):
    """uint32 [..., 4] -> (bool [..., 88] key activations, bool [...] pedal)."""
    w = np.asarray(words, np.uint32)
    bits = ((w[..., :3, None] >> np.arange(32, dtype=np.uint32)) & 1).astype(bool)
    return bits.reshape(*w.shape[:-1], 96)[..., :NUM_KEYS], w[..., 3] != 0


class MidiModule:
    def __init__(self) -> None:
        self._note_on_callback: Optional[Callable[[int, int], None]] = None
        self._note_off_callback: Optional[Callable[[int], None]] = None
        self._sustain_on_callback: Optional[Callable[[], None]] = None
        self._sustain_off_callback: Optional[Callable[[], None]] = None
        self.initialize_episode()

    def initialize_episode(self, physics=None) -> None:
        del physics
        self._prev_activation = np.zeros(NUM_KEYS, dtype=bool)
        self._prev_sustain_activation = False
        self._midi_messages: List[List[MidiMessage]] = []

    def after_substep(self, time: float, activation: np.ndarray, sustain_activation: bool) -> None:
        """midi_module.py:47-98, with the substep's simulation time passed in instead of `physics`."""
        activation = np.asarray(activation, dtype=bool)
        sustain_activation = bool(sustain_activation)
        events: List[MidiMessage] = []
        change = activation ^ self._prev_activation
        for key in np.flatnonzero(change & ~self._prev_activation):
            msg = NoteOn(note=int(key) + MIN_MIDI_PITCH_PIANO, velocity=127, time=time)
            events.append(msg)
            if self._note_on_callback is not None:
                self._note_on_callback(msg.note, msg.velocity)
        for key in np.flatnonzero(change & ~activation):
            msg = NoteOff(note=int(key) + MIN_MIDI_PITCH_PIANO, time=time)
            events.append(msg)
            if self._note_off_callback is not None:
                self._note_off_callback(msg.note)
        if sustain_activation and not self._prev_sustain_activation:
            events.append(SustainOn(time=time))
            if self._sustain_on_callback is not None:
                self._sustain_on_callback()
        if self._prev_sustain_activation and not sustain_activation:
            events.append(SustainOff(time=time))
            if self._sustain_off_callback is not None:
                self._sustain_off_callback()
        self._midi_messages.append(events)
        self._prev_activation = activation.copy()
        self._prev_sustain_activation = sustain_activation

    def after_step_words(self, words: np.ndarray, first_substep_index: int, timestep: float) -> None:
        """One control step's recording for one env: uint32 [n_substeps, 4]; substep j ends at simulation time
        (first_substep_index + j + 1) * timestep."""
        act, sus = unpack_activation_words(words)
        for j in range(act.shape[0]):
            self.after_substep((first_substep_index + j + 1) * timestep, act[j], sus[j])

    def get_latest_midi_messages(self) -> List[MidiMessage]:
        return self._midi_messages[-1]

    def get_all_midi_messages(self) -> List[MidiMessage]:
        return [m for step in self._midi_messages for m in step]

    def register_synth_note_on_callback(self, callback: Callable[[int, int], None]) -> None:
        self._note_on_callback = callback

    def register_synth_note_off_callback(self, callback: Callable[[int], None]) -> None:
        self._note_off_callback = callback

    def register_synth_sustain_on_callback(self, callback: Callable[[], None]) -> None:
        self._sustain_on_callback = callback

    def register_synth_sustain_off_callback(self, callback: Callable[[], None]) -> None:
        self._sustain_off_callback = callback
