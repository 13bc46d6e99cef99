"""Song library: the reference's programmatic debug songs plus the bundled MIDI cuts.

Same names and note content as reference `robopianist/music/library.py` (toy :28-67,
one/two-hand scales :69-300, chord progression :303-334, Rousseau cuts :337-396 + nocturne,
MIDI_NAME_TO_CALLABLE :544-553). The MIDI-backed songs are stored as tick-level event
lists in data/songs.json (see tools/make_song_fixtures.py) and go through this package's
own SMF->NoteSequence rules at load time.
"""

from __future__ import annotations

import json
from functools import lru_cache
from pathlib import Path
from typing import Callable, Dict

from . import midi_file
from .sequence import NoteSequence
from .smf import events_to_sequence

_DATA = Path(__file__).resolve().parent.parent / "data"


@lru_cache(maxsize=1)
def _songs() -> dict:
    return json.loads((_DATA / "songs.json").read_text())


def _seq(title: str, total_time: float) -> NoteSequence:
    seq = NoteSequence()
    seq.sequence_metadata.title = title
    seq.sequence_metadata.artist = "robopianist"
    seq.total_time = total_time
    seq.tempos.add(qpm=60)
    return seq


def toy(right_finger: int = 1, left_finger: int = 6) -> midi_file.MidiFile:
    seq = NoteSequence()
    n2m = midi_file.note_name_to_midi_number
    for name, t0, part in (("C6", 0.0, right_finger), ("G5", 0.5, right_finger),
                           ("C3", 0.0, left_finger), ("C4", 0.5, left_finger)):
        seq.notes.add(start_time=t0, end_time=t0 + 0.5, velocity=80, pitch=n2m(name), part=part)
    seq.total_time = 1.0
    seq.tempos.add(qpm=60)
    return midi_file.MidiFile(seq=seq)


def twinkle_twinkle_little_star_one_hand() -> midi_file.MidiFile:
    seq = _seq("Twinkle Twinkle (one hand)", 8.0)
    rows = [(60, 0.0, 0.5, 0), (60, 0.5, 1.0, 0), (67, 1.0, 1.5, 2), (67, 1.5, 2.0, 2),
            (69, 2.0, 2.5, 3), (69, 2.5, 3.0, 3), (67, 3.0, 4.0, 2), (65, 4.0, 4.5, 3),
            (65, 4.5, 5.0, 3), (64, 5.0, 5.5, 2), (64, 5.5, 6.0, 2), (62, 6.0, 6.5, 1),
            (62, 6.5, 7.0, 1), (60, 7.0, 8.0, 0)]
    for pitch, s, e, part in rows:
        seq.notes.add(pitch=pitch, start_time=s, end_time=e, velocity=80, part=part)
    return midi_file.MidiFile(seq=seq)


_C_MAJOR = [0, 2, 4, 5, 7, 9, 11, 12]
_D_MAJOR = [2, 4, 6, 7, 9, 11, 13, 14]
_RH_UP, _RH_DOWN = [0, 1, 2, 0, 1, 2, 3, 4], [3, 2, 1, 0, 2, 1, 0]
_LH_UP, _LH_DOWN = [9, 8, 7, 6, 5, 7, 6, 5], [6, 7, 5, 6, 7, 8, 9]


def _scale(title, degrees, right_octave, left_octave, dur) -> midi_file.MidiFile:
    seq = _seq(title, 15 * dur)
    for i in range(15):
        up = i < 8
        deg = degrees[i] if up else degrees[7 - (i - 8) - 1]
        seq.notes.add(pitch=12 * right_octave + deg, start_time=i * dur, end_time=(i + 1) * dur,
                      velocity=80, part=_RH_UP[i] if up else _RH_DOWN[i - 8])
        if left_octave is not None:
            seq.notes.add(pitch=12 * left_octave + deg, start_time=i * dur, end_time=(i + 1) * dur,
                          velocity=80, part=_LH_UP[i] if up else _LH_DOWN[i - 8])
    return midi_file.MidiFile(seq=seq)


def c_major_scale_one_hand(right_octave: int = 6, note_duration: float = 0.5):
    return _scale("C major scale (one hand)", _C_MAJOR, right_octave, None, note_duration)


def d_major_scale_one_hand(right_octave: int = 6, note_duration: float = 0.5):
    return _scale("D major scale (one hand)", _D_MAJOR, right_octave, None, note_duration)


def c_major_scale_two_hands(left_octave: int = 4, right_octave: int = 6, note_duration: float = 0.5):
    return _scale("C major scale", _C_MAJOR, right_octave, left_octave, note_duration)


def d_major_scale_two_hands(left_octave: int = 4, right_octave: int = 6, note_duration: float = 0.5):
    return _scale("D major scale", _D_MAJOR, right_octave, left_octave, note_duration)


def c_major_chord_progression_two_hands() -> midi_file.MidiFile:
    seq = _seq("C major chord progression", 4)
    chords = [((48, 5), (60, 0), (64, 2), (67, 4)), ((41, 8), (65, 0), (69, 2), (72, 4)),
              ((43, 7), (67, 0), (71, 2), (74, 4)), ((48, 5), (60, 0), (64, 2), (67, 4))]
    for t, chord in enumerate(chords):
        for pitch, part in chord:
            seq.notes.add(pitch=pitch, start_time=t, end_time=t + 1, velocity=80, part=part)
    return midi_file.MidiFile(seq=seq)


def _bundled(name: str) -> midi_file.MidiFile:
    entry = _songs()[name]
    tracks = [[tuple(e) for e in t] for t in entry["tracks"]]
    midi = midi_file.MidiFile(seq=events_to_sequence(entry["resolution"], tracks))
    if "title" in entry:
        midi.seq.sequence_metadata.artist = entry["artist"]
        midi.seq.sequence_metadata.title = entry["title"]
    if "fingering" in entry:
        # fingering listed in start-time order (stable), reference library.py:387-394
        ordered = sorted(midi.seq.notes, key=lambda n: n.start_time)
        assert len(ordered) == len(entry["fingering"])
        for note, finger in zip(ordered, entry["fingering"]):
            note.part = finger
    return midi


def twinkle_twinkle_rousseau() -> midi_file.MidiFile:
    return _bundled("twinkle_twinkle_rousseau")


def nocturne_rousseau() -> midi_file.MidiFile:
    return _bundled("nocturne_rousseau")


def guren_no_yumiya_cut_14s(with_fingering: bool = True) -> midi_file.MidiFile:
    """Config 3 song: reference midi_files_cut/ + its fingering annotation file, merged
    by the rule of data_processing/add_fingering_to_midi.py:66-75."""
    from .fingering import apply_annotation_rows

    midi = _bundled("guren_no_yumiya_cut_14s")
    if with_fingering:
        apply_annotation_rows(midi.seq, _songs()["guren_no_yumiya_cut_14s"]["annotation"])
    return midi


def crossing_field_cut_10s(with_fingering: bool = True) -> midi_file.MidiFile:
    """Config 4 song. Its annotation file is in ticks, so no note matches and the song
    ends up without fingering - reproduced, not fixed."""
    from .fingering import apply_annotation_rows

    midi = _bundled("crossing_field_cut_10s")
    if with_fingering:
        apply_annotation_rows(midi.seq, _songs()["crossing_field_cut_10s"]["annotation"])
    return midi


MIDI_NAME_TO_CALLABLE: Dict[str, Callable[[], midi_file.MidiFile]] = {
    "TwinkleTwinkleLittleStar": twinkle_twinkle_little_star_one_hand,
    "CMajorScaleOneHand": c_major_scale_one_hand,
    "CMajorScaleTwoHands": c_major_scale_two_hands,
    "DMajorScaleOneHand": d_major_scale_one_hand,
    "DMajorScaleTwoHands": d_major_scale_two_hands,
    "CMajorChordProgressionTwoHands": c_major_chord_progression_two_hands,
    "TwinkleTwinkleRousseau": twinkle_twinkle_rousseau,
    "NocturneRousseau": nocturne_rousseau,
    # bundled cuts the reference's training scripts load by absolute path
    "GurenNoYumiyaCut14s": guren_no_yumiya_cut_14s,
    "CrossingFieldCut10s": crossing_field_cut_10s,
}
