"""MIDI augmentations (reference robopianist/suite/variations.py:27-184): the cases of the reference's
variations_test.py ported against suite/variations.py, plus the draw order on a seeded RandomState."""
import numpy as np
import pytest

from diffusion_piano_b200 import music, suite
from diffusion_piano_b200.music import midi_file
from diffusion_piano_b200.suite import variations

_SEED = 12345
_NUM_SAMPLES = 100


def _toy():
    return music.load("CMajorScaleTwoHands")


def test_midi_select_output_is_midi_file():
    var = variations.MidiSelect(midi_names=["CMajorScaleOneHand", "CMajorScaleTwoHands", "TwinkleTwinkleLittleStar"])
    rs = np.random.RandomState(_SEED)
    titles = set()
    for _ in range(20):
        midi = var(random_state=rs)
        assert isinstance(midi, midi_file.MidiFile)
        titles.add(len(midi.seq.notes))
    assert len(titles) > 1  # it does select different files


@pytest.mark.parametrize("make", [lambda p: variations.MidiTemporalStretch(prob=p, stretch_range=0.5),
                                  lambda p: variations.MidiPitchShift(prob=p, shift_range=1),
                                  lambda p: variations.MidiOctaveShift(prob=p, octave_range=1)])
def test_probability_semantics(make):
    original = _toy()
    rs = np.random.RandomState(_SEED)
    for _ in range(_NUM_SAMPLES):  # output is a MidiFile
        assert isinstance(make(0.1)(initial_value=original, random_state=rs), midi_file.MidiFile)
    for _ in range(_NUM_SAMPLES):  # prob 0: the very same object comes back
        assert make(0.0)(initial_value=original, random_state=rs) is original
    with pytest.raises(ValueError):  # not applied and nothing to hand back
        make(0.0)(random_state=rs)
    with pytest.raises(ValueError):
        make(0.0)(initial_value=1, random_state=rs)


def test_stretch_prob_one_changes_the_file_and_range_zero_does_not():
    original = _toy()
    rs = np.random.RandomState(_SEED)
    var = variations.MidiTemporalStretch(prob=1.0, stretch_range=0.5)
    for _ in range(20):
        new = var(initial_value=original, random_state=rs)
        assert new is not original and new.seq.total_time != original.seq.total_time
        assert 0.5 * original.seq.total_time <= new.seq.total_time <= 1.5 * original.seq.total_time
    same = variations.MidiTemporalStretch(prob=1.0, stretch_range=0.0)(initial_value=original, random_state=rs)
    assert same.seq.total_time == original.seq.total_time
    assert [(n.pitch, n.start_time, n.end_time) for n in same.seq.notes] == \
           [(n.pitch, n.start_time, n.end_time) for n in original.seq.notes]


def test_pitch_and_octave_shift_stay_on_the_piano():
    original = _toy()
    rs = np.random.RandomState(_SEED)
    assert variations.MidiPitchShift(prob=1.0, shift_range=0)(initial_value=original, random_state=rs) is original
    assert variations.MidiOctaveShift(prob=1.0, octave_range=0)(initial_value=original, random_state=rs) is original
    base = [n.pitch for n in original.seq.notes]
    seen = set()
    for _ in range(50):
        new = variations.MidiPitchShift(prob=1.0, shift_range=40)(initial_value=original, random_state=rs)
        d = {n.pitch - b for n, b in zip(new.seq.notes, base)}
        assert len(d) == 1 and all(21 <= n.pitch <= 108 for n in new.seq.notes)
        seen |= d
    assert len(seen) > 5 and min(seen) >= 21 - min(base) and max(seen) <= 108 - max(base)  # truncated to the piano
    # octave_range 2 keeps C3..C6 on the piano. (With 3 the reference's `low // 12` floors -27 semitones to -3
    # octaves and the lowest notes are deleted by transpose - variations.py:176-181; the mirror keeps that quirk.)
    for _ in range(20):
        new = variations.MidiOctaveShift(prob=1.0, octave_range=2)(initial_value=original, random_state=rs)
        d = {n.pitch - b for n, b in zip(new.seq.notes, base)}
        assert len(d) == 1 and next(iter(d)) % 12 == 0 and all(21 <= n.pitch <= 108 for n in new.seq.notes)
    with pytest.raises(ValueError):
        variations.MidiPitchShift(prob=0.5, shift_range=1.5)
    with pytest.raises(ValueError):
        variations.MidiOctaveShift(prob=0.5, octave_range=0.5)


def test_draw_order_on_a_seeded_state():
    """The reference draws the probability first, then the value (variations.py:68-75): restated on numpy."""
    original = _toy()
    rs, rs2 = np.random.RandomState(3), np.random.RandomState(3)
    new = variations.MidiTemporalStretch(prob=0.7, stretch_range=0.3)(initial_value=original, random_state=rs)
    u = rs2.uniform(0.0, 1.0)
    if u > 0.7:
        assert new is original
    else:
        factor = 1.0 + rs2.uniform(-0.3, 0.3)
        np.testing.assert_allclose(new.seq.total_time, original.seq.total_time * factor, rtol=1e-12)


def test_task_draws_songs_through_its_augmentations():
    """piano_with_shadow_hands.py:151-165: the augmentations are applied in order to the INITIAL midi."""
    task = suite.PianoWithShadowHands(_toy(), augmentations=[variations.MidiTemporalStretch(1.0, 0.4),
                                                            variations.MidiPitchShift(1.0, 5)])
    rs = np.random.RandomState(0)
    frames = set()
    for _ in range(8):
        midi, notes, sustains, tables = task.draw_song(rs)
        assert len(notes) == len(sustains) == tables.n_frames == tables.goal.shape[0]
        frames.add(tables.n_frames)
    assert len(frames) > 2
    assert task.song.n_frames == len(task._notes)  # the task's own trajectory is still the initial midi's
