"""Music constants (same values as reference robopianist/music/constants.py)."""

MIN_MIDI_PITCH = 0
MAX_MIDI_PITCH = 127
MIN_MIDI_PITCH_PIANO = 21  # A0
MAX_MIDI_PITCH_PIANO = 108  # C8
MIN_KEY_NUMBER = 0
MAX_KEY_NUMBER = 87
NUM_KEYS = 88

_OCTAVE = ("C", "C#", "D", "D#", "E", "F", "F#", "G", "G#", "A", "A#", "B")
NOTES = ["A0", "A#0", "B0"] + [n + str(o) for o in range(1, 8) for n in _OCTAVE] + ["C8"]
assert len(NOTES) == NUM_KEYS

KEY_NUMBER_TO_NOTE_NAME = dict(enumerate(NOTES))
NOTE_NAME_TO_KEY_NUMBER = {n: i for i, n in enumerate(NOTES)}
MIDI_NUMBER_TO_NOTE_NAME = {i + MIN_MIDI_PITCH_PIANO: n for i, n in enumerate(NOTES)}
NOTE_NAME_TO_MIDI_NUMBER = {n: i for i, n in MIDI_NUMBER_TO_NOTE_NAME.items()}

SAMPLING_RATE = 44100
SUSTAIN_PEDAL_CC_NUMBER = 64
MIN_CC_VALUE = 0
MAX_CC_VALUE = 127
MIN_VELOCITY = 0
MAX_VELOCITY = 127
