"""A minimal, dependency-free stand-in for the `note_seq.NoteSequence` container.

The reference stores songs as `note_seq` protobufs (reference
`robopianist/music/midi_file.py:30,166-237`). note_seq is a third-party package that
is not vendored in the reference tree, so this module restates the parts of its
published behaviour the hot path's goal tables depend on:

* the container itself (notes, control changes, tempos, total_time, metadata);
* `sequences_lib.stretch_note_sequence` (used by `MidiFile.stretch`, midi_file.py:199-209);
* `sequences_lib.transpose_note_sequence` (`MidiFile.transpose`, midi_file.py:211-229);
* `sequences_lib.extract_subsequence` (`MidiFile.trim_silence`, midi_file.py:231-237).

Fingering is carried in `Note.part`, exactly like the reference (protobuf default 0).
"""

from __future__ import annotations

import copy
from dataclasses import dataclass, field
from typing import List

SUSTAIN_CONTROL_NUMBER = 64


@dataclass
class Note:
    pitch: int = 0
    velocity: int = 0
    start_time: float = 0.0
    end_time: float = 0.0
    part: int = 0  # fingering lives here (protobuf int default is 0, never None)
    instrument: int = 0
    program: int = 0
    is_drum: bool = False


@dataclass
class ControlChange:
    time: float = 0.0
    control_number: int = 0
    control_value: int = 0
    instrument: int = 0
    program: int = 0
    is_drum: bool = False


@dataclass
class Tempo:
    time: float = 0.0
    qpm: float = 120.0


@dataclass
class TimeSignature:
    time: float = 0.0
    numerator: int = 4
    denominator: int = 4


@dataclass
class SequenceMetadata:
    title: str = ""
    artist: str = ""


class _Repeated(list):
    """list with a protobuf-style `.add(**kw)` so reference-style construction reads the same."""

    def __init__(self, cls, *a):
        super().__init__(*a)
        self._cls = cls

    def add(self, **kw):
        item = self._cls(**kw)
        self.append(item)
        return item

    def __deepcopy__(self, memo):
        return _Repeated(self._cls, [copy.deepcopy(x, memo) for x in self])


@dataclass
class NoteSequence:
    notes: List[Note] = field(default_factory=lambda: _Repeated(Note))
    control_changes: List[ControlChange] = field(
        default_factory=lambda: _Repeated(ControlChange)
    )
    tempos: List[Tempo] = field(default_factory=lambda: _Repeated(Tempo))
    time_signatures: List[TimeSignature] = field(
        default_factory=lambda: _Repeated(TimeSignature)
    )
    total_time: float = 0.0
    ticks_per_quarter: int = 220
    sequence_metadata: SequenceMetadata = field(default_factory=SequenceMetadata)

    def copy(self) -> "NoteSequence":
        return copy.deepcopy(self)


def stretch_note_sequence(seq: NoteSequence, stretch_factor: float) -> NoteSequence:
    """note_seq `sequences_lib.stretch_note_sequence`: scale every time by the factor.

    factor == 1 returns an unmodified copy (reference test midi_file_test.py:41-44).
    """
    out = seq.copy()
    if stretch_factor == 1.0:
        return out
    for n in out.notes:
        n.start_time *= stretch_factor
        n.end_time *= stretch_factor
    out.total_time *= stretch_factor
    for cc in out.control_changes:
        cc.time *= stretch_factor
    for ts in out.time_signatures:
        ts.time *= stretch_factor
    for t in out.tempos:
        t.time *= stretch_factor
        t.qpm /= stretch_factor
    return out


def transpose_note_sequence(
    seq: NoteSequence, amount: int, min_allowed_pitch: int, max_allowed_pitch: int
):
    """note_seq `sequences_lib.transpose_note_sequence` (non-drum notes shifted,
    out-of-range notes deleted). Returns (sequence, num_deleted)."""
    out = seq.copy()
    kept = _Repeated(Note)
    deleted = 0
    for n in out.notes:
        if not n.is_drum:
            n.pitch += amount
            if n.pitch < min_allowed_pitch or n.pitch > max_allowed_pitch:
                deleted += 1
                continue
        kept.append(n)
    out.notes = kept
    if deleted:
        out.total_time = max((n.end_time for n in out.notes), default=0.0)
    return out, deleted


def extract_subsequence(
    seq: NoteSequence, start_time: float, end_time: float
) -> NoteSequence:
    """note_seq `sequences_lib.extract_subsequence`.

    Notes whose *start* lies in [start_time, end_time) are kept, shifted by
    -start_time and truncated at end_time. Only sustain (CC64) control changes
    survive; the last sustain event at or before start_time is re-emitted per
    instrument at t=0, appended *after* the in-range events.
    """
    if start_time >= end_time:
        raise ValueError("start_time must be < end_time")
    if start_time >= seq.total_time:
        raise ValueError("start_time must be < sequence.total_time")
    out = seq.copy()
    out.notes = _Repeated(Note)
    for n in seq.notes:
        if n.start_time < start_time or n.start_time >= end_time:
            continue
        m = copy.copy(n)
        m.start_time = n.start_time - start_time
        m.end_time = min(n.end_time, end_time) - start_time
        out.notes.append(m)
    out.total_time = min(seq.total_time, end_time) - start_time

    # tempos / time signatures: keep the state in force at start_time at t=0, plus
    # in-range events.
    for name in ("tempos", "time_signatures"):
        events = sorted(getattr(seq, name), key=lambda e: e.time)
        kept = _Repeated(type(events[0])) if events else _Repeated(Tempo)
        initial = None
        for e in events:
            if e.time <= start_time:
                initial = e
            elif e.time < end_time:
                c = copy.copy(e)
                c.time = e.time - start_time
                kept.append(c)
        if initial is not None:
            c = copy.copy(initial)
            c.time = 0.0
            kept.insert(0, c)
        setattr(out, name, kept)

    out.control_changes = _Repeated(ControlChange)
    sustain = sorted(
        (c for c in seq.control_changes if c.control_number == SUSTAIN_CONTROL_NUMBER),
        key=lambda c: c.time,
    )
    initial = {}
    for c in sustain:
        if c.time <= start_time:
            initial[c.instrument] = c
            continue
        if c.time >= end_time:
            break
        m = copy.copy(c)
        m.time = c.time - start_time
        out.control_changes.append(m)
    for _, c in sorted(initial.items()):
        m = copy.copy(c)
        m.time = 0.0
        out.control_changes.append(m)
    return out
