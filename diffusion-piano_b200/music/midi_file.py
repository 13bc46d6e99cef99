"""MidiFile / PianoNote / NoteTrajectory: the song -> per-control-step goal tables.

Host-side, once per song. Mirrors the public surface of the
reference's `robopianist/music/midi_file.py` (PianoNote :116-162, MidiFile :165-282,
NoteTrajectory :285-414) and the frame quantisation of
`robopianist/music/piano_roll.py:59-204`, without note_seq/pretty_midi.

The dense device tables (`SongTables`) are what the CUDA step consumes.
"""

from __future__ import annotations

import math
from dataclasses import dataclass
from pathlib import Path
from typing import List, Tuple, Union

import numpy as np

from . import constants as consts
from . import sequence as seqlib
from .sequence import NoteSequence
from .smf import MidiParseError, read_midi, write_smf

MIN_MIDI_VELOCITY = 1  # note_seq.constants.MIN_MIDI_VELOCITY
MAX_MIDI_VELOCITY = 127


def note_name_to_midi_number(name: str) -> int:
    return consts.NOTE_NAME_TO_MIDI_NUMBER[name]


def midi_number_to_note_name(number: int) -> str:
    return consts.MIDI_NUMBER_TO_NOTE_NAME[number]


def key_number_to_midi_number(key_number: int) -> int:
    if not 0 <= key_number < consts.NUM_KEYS:
        raise ValueError(f"Key number should be in [0 {consts.NUM_KEYS}], got {key_number}.")
    return key_number + consts.MIN_MIDI_PITCH_PIANO


def midi_number_to_key_number(midi_number: int) -> int:
    if not consts.MIN_MIDI_PITCH_PIANO <= midi_number <= consts.MAX_MIDI_PITCH_PIANO:
        raise ValueError(
            f"MIDI pitch number should be in [{consts.MIN_MIDI_PITCH_PIANO}, "
            f"{consts.MAX_MIDI_PITCH_PIANO}], got {midi_number}."
        )
    return midi_number - consts.MIN_MIDI_PITCH_PIANO


def key_number_to_note_name(key_number: int) -> str:
    return consts.KEY_NUMBER_TO_NOTE_NAME[key_number]


def note_name_to_key_number(note_name: str) -> int:
    return consts.NOTE_NAME_TO_KEY_NUMBER[note_name]


@dataclass(frozen=True)
class PianoNote:
    """reference midi_file.py:116-162. fingering: 0-4 right thumb..pinky, 5-9 left."""

    number: int
    velocity: int
    key: int
    name: str
    fingering: int = -1

    @staticmethod
    def create(number: int, velocity: int, fingering: int = -1) -> "PianoNote":
        if not MIN_MIDI_VELOCITY <= velocity <= MAX_MIDI_VELOCITY:
            raise ValueError(
                f"Velocity should be in [{MIN_MIDI_VELOCITY}, {MAX_MIDI_VELOCITY}], got {velocity}."
            )
        key = midi_number_to_key_number(number)
        return PianoNote(number, velocity, key, midi_number_to_note_name(number), fingering)


@dataclass(frozen=True)
class MidiFile:
    """reference midi_file.py:165-282 (audio synthesis / playback are out of scope)."""

    seq: NoteSequence

    @classmethod
    def from_file(cls, filename: Union[str, Path]) -> "MidiFile":
        filename = Path(filename)
        if filename.suffix != ".mid":
            raise ValueError(f"Unsupported file extension {filename.suffix}.")
        try:
            return cls(seq=read_midi(filename))
        except MidiParseError as e:
            raise RuntimeError(f"Could not parse MIDI file {filename}.") from e

    def save(self, filename: Union[str, Path]) -> None:
        filename = Path(filename)
        if filename.suffix != ".mid":
            raise ValueError(f"Unsupported file extension {filename.suffix}.")
        filename.write_bytes(write_smf(self.seq))

    def stretch(self, factor: float) -> "MidiFile":
        if factor <= 0:
            raise ValueError("factor must be positive.")
        return MidiFile(seq=seqlib.stretch_note_sequence(self.seq, factor))

    def transpose(self, amount: int, transpose_chords: bool = True) -> "MidiFile":
        del transpose_chords
        seq, _ = seqlib.transpose_note_sequence(
            self.seq, amount, consts.MIN_MIDI_PITCH_PIANO, consts.MAX_MIDI_PITCH_PIANO
        )
        return MidiFile(seq=seq)

    def trim_silence(self) -> "MidiFile":
        # list order on purpose (reference midi_file.py:231-237)
        return MidiFile(
            seq=seqlib.extract_subsequence(
                self.seq, self.seq.notes[0].start_time, self.seq.notes[-1].end_time
            )
        )

    def has_fingering(self) -> bool:
        parts = {n.part for n in self.seq.notes}
        return len(parts) > 1 and any(p != 0 for p in parts)

    @property
    def duration(self) -> float:
        return self.seq.total_time

    @property
    def n_notes(self) -> int:
        return len(self.seq.notes)

    @property
    def title(self) -> str:
        return self.seq.sequence_metadata.title

    @property
    def artist(self) -> str:
        return self.seq.sequence_metadata.artist


def _frame_span(start: float, end: float, fps: float) -> Tuple[int, int]:
    """piano_roll.py:94-117 with min_frame_occupancy_for_label == 0."""
    a = int(start * fps)
    b = int(math.ceil(end * fps))
    return a, max(a + 1, b)


@dataclass
class PianoRoll:
    active_velocities: np.ndarray  # (T,128) float32, velocity/127 where a note sounds
    onset_velocities: np.ndarray  # (T,128) float32, velocity on the onset frame only
    fingerings: np.ndarray  # (T,128) float32, -1 where nothing wrote
    control_changes: np.ndarray  # (T,128) int32, value+1 on the event frame


def sequence_to_pianoroll(seq: NoteSequence, frames_per_second: float) -> PianoRoll:
    """The subset of piano_roll.py:59-204 that NoteTrajectory consumes
    (min_pitch=0, max_pitch=127, onset_window=0, onset_mode='window')."""
    n_frames = int(seq.total_time * frames_per_second + 1)
    vel = np.zeros((n_frames, 128), np.float32)
    onset = np.zeros((n_frames, 128), np.float32)
    fing = np.full((n_frames, 128), -1.0, np.float32)
    cc = np.zeros((n_frames, 128), np.int32)
    # stable sort by start time: later-starting notes overwrite earlier ones
    for note in sorted(seq.notes, key=lambda n: n.start_time):
        if not 0 <= note.pitch <= 127:
            continue
        a, b = _frame_span(note.start_time, note.end_time, frames_per_second)
        if note.velocity > MAX_MIDI_VELOCITY:
            raise ValueError(
                "Note velocity exceeds max velocity: %d > %d" % (note.velocity, MAX_MIDI_VELOCITY)
            )
        on_a = max(0, a)
        on_b = min(n_frames, a + 1)
        onset[on_a:on_b, note.pitch] = 1.0
        vel[a:b, note.pitch] = note.velocity / MAX_MIDI_VELOCITY
        if note.part is not None:
            fing[a:b, note.pitch] = note.part
    for c in seq.control_changes:
        frame = int(c.time * frames_per_second)
        if frame < n_frames:
            cc[frame, c.control_number] = c.control_value + 1
    return PianoRoll(vel, vel * onset, fing, cc)


@dataclass
class NoteTrajectory:
    """reference midi_file.py:285-414."""

    dt: float
    notes: List[List[PianoNote]]
    sustains: List[int]

    def __post_init__(self) -> None:
        if self.dt <= 0:
            raise ValueError("dt must be positive.")
        if len(self.notes) != len(self.sustains):
            raise ValueError("notes and sustains must have the same length.")

    @classmethod
    def from_midi(cls, midi: MidiFile, dt: float) -> "NoteTrajectory":
        notes, sustains = cls.seq_to_trajectory(midi.seq, dt)
        return cls(dt=dt, notes=notes, sustains=sustains)

    @staticmethod
    def seq_to_trajectory(seq: NoteSequence, dt: float):
        roll = sequence_to_pianoroll(seq, frames_per_second=1 / dt)
        av, ov = roll.active_velocities, roll.onset_velocities
        notes: List[List[PianoNote]] = []
        for t in range(av.shape[0]):
            frame: List[PianoNote] = []
            for pitch in np.nonzero(av[t])[0]:
                # re-struck key: leave one empty frame so the key is released first
                if t > 0 and av[t - 1][pitch] and ov[t][pitch]:
                    continue
                velocity = int(round(av[t][pitch] * consts.MAX_VELOCITY))
                frame.append(PianoNote.create(int(pitch), velocity, int(roll.fingerings[t, pitch])))
            notes.append(frame)
        sustains: List[int] = []
        prev = 0
        for t in range(roll.control_changes.shape[0]):
            ev = roll.control_changes[t][consts.SUSTAIN_PEDAL_CC_NUMBER]
            if 1 <= ev <= consts.SUSTAIN_PEDAL_CC_NUMBER:
                cur = 0
            elif consts.SUSTAIN_PEDAL_CC_NUMBER + 1 <= ev <= consts.MAX_CC_VALUE + 1:
                cur = 1
            else:
                cur = prev
            sustains.append(cur)
            prev = cur
        return notes, sustains

    def __len__(self) -> int:
        return len(self.notes)

    def add_initial_buffer_time(self, initial_buffer_time: float) -> "NoteTrajectory":
        if initial_buffer_time < 0.0:
            raise ValueError("initial_buffer_time must be non-negative.")
        for _ in range(int(round(initial_buffer_time / self.dt))):
            self.notes.insert(0, [])
            self.sustains.insert(0, 0)
        return self

    def to_piano_roll(self) -> np.ndarray:
        frames = np.zeros((len(self.notes), consts.MAX_MIDI_PITCH), dtype=np.int32)
        for t, step in enumerate(self.notes):
            for note in step:
                frames[t, note.number] = 1
        return frames


@dataclass
class SongTables:
    """Dense per-song tables uploaded once to HBM (shared by all envs on a GPU).

    goal[t, 0:88] = 1 where key is to be held in frame t, goal[t, 88] = sustain.
    pairs[t, p] = (key, finger 0..9) for the p-th note of frame t in reference order
    (np.nonzero order = ascending key), key = -1 padding. A finger value outside 0..9
    (only possible when a test sets part=-1) is stored as-is.
    """

    goal: np.ndarray  # (T, 89) uint8
    pairs: np.ndarray  # (T, pmax, 2) int16
    n_frames: int
    pmax: int

    @classmethod
    def from_trajectory(cls, traj: NoteTrajectory, min_pmax: int = 1) -> "SongTables":
        T = len(traj.notes)
        pmax = max([len(f) for f in traj.notes] + [min_pmax])
        goal = np.zeros((T, consts.NUM_KEYS + 1), np.uint8)
        pairs = np.full((T, pmax, 2), -1, np.int16)
        for t, frame in enumerate(traj.notes):
            for p, note in enumerate(frame):
                goal[t, note.key] = 1
                pairs[t, p, 0] = note.key
                pairs[t, p, 1] = note.fingering
            goal[t, consts.NUM_KEYS] = traj.sustains[t]
        return cls(goal=goal, pairs=pairs, n_frames=T, pmax=pmax)
