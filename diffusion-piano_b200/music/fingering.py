"""Fingering annotation loader.

Mirrors reference `data_processing/add_fingering_to_midi.py:26-83`: 8-column
tab-separated rows `id start end pitch vel vel hand finger`; a note receives
`part = finger` from the first row with |start diff| < 0.01, |end diff| < 0.01 and equal
pitch; finger values outside 0..9 are ignored.
"""

from __future__ import annotations

import re
from pathlib import Path
from typing import Iterable, Sequence, Union

from .midi_file import MidiFile
from .sequence import NoteSequence

_NOTE_VALUES = {"C": 0, "C#": 1, "Db": 1, "D": 2, "D#": 3, "Eb": 3, "E": 4, "F": 5, "F#": 6,
                "Gb": 6, "G": 7, "G#": 8, "Ab": 8, "A": 9, "A#": 10, "Bb": 10, "B": 11}


def parse_pitch_to_midi_number(pitch_str: str) -> int:
    m = re.match(r"([A-G][#b]?)(\d+)", pitch_str)
    if not m:
        raise ValueError(f"Invalid pitch format: {pitch_str}")
    note, octave = m.groups()
    return _NOTE_VALUES[note] + (int(octave) + 1) * 12


def apply_annotation_rows(seq: NoteSequence, rows: Iterable[Sequence[str]]) -> int:
    table = []
    for parts in rows:
        if len(parts) != 8:
            continue
        _, start, end, pitch, _, _, _hand, finger = parts
        f = int(finger)
        if 0 <= f <= 9:
            table.append((float(start), float(end), parse_pitch_to_midi_number(pitch), f))
    matched = 0
    for note in seq.notes:
        for start, end, pitch, f in table:
            if abs(note.start_time - start) < 0.01 and abs(note.end_time - end) < 0.01 and note.pitch == pitch:
                note.part = f
                matched += 1
                break
    return matched


def read_annotation_rows(path: Union[str, Path]):
    rows = []
    for line in Path(path).read_text().splitlines():
        if line.startswith("//") or not line.strip():
            continue
        rows.append(line.strip().split("\t"))
    return rows


def add_fingering_from_annotation_file(input_midi: Union[str, Path], annotation_file: Union[str, Path]) -> MidiFile:
    midi = MidiFile.from_file(Path(input_midi))
    apply_annotation_rows(midi.seq, read_annotation_rows(annotation_file))
    return midi
