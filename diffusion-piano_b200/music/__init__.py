"""Music module: `load(name_or_path, stretch, shift) -> MidiFile`
(reference robopianist/music/__init__.py:57-92; the PIG dataset is not bundled there
either, so PIG_MIDIS is empty unless files are supplied)."""

from pathlib import Path
from typing import Union

from . import library, midi_file
from .midi_file import MidiFile, NoteTrajectory, PianoNote, SongTables

DEBUG_MIDIS = list(library.MIDI_NAME_TO_CALLABLE.keys())
PIG_MIDIS: list = []
ETUDE_MIDIS: list = []
ALL = DEBUG_MIDIS + PIG_MIDIS


def load(path_or_name: Union[str, Path], stretch: float = 1.0, shift: int = 0) -> MidiFile:
    path = Path(path_or_name)
    if path.suffix:
        midi = MidiFile.from_file(path)
    elif path.stem in DEBUG_MIDIS:
        midi = library.MIDI_NAME_TO_CALLABLE[path.stem]()
    else:
        raise KeyError(f"Unknown name: {path.stem}. Available names: {ALL}.")
    return midi.stretch(stretch).transpose(shift)


__all__ = ["ALL", "DEBUG_MIDIS", "PIG_MIDIS", "ETUDE_MIDIS", "load", "MidiFile",
           "NoteTrajectory", "PianoNote", "SongTables"]
