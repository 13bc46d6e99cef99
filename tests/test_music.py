"""Host music layer: ported from reference robopianist/music/midi_file_test.py (:30-215) and
checked against numbers read off the reference's data."""
import numpy as np
import pytest

from diffusion_piano_b200 import music
from diffusion_piano_b200.music import midi_file
from diffusion_piano_b200.music.sequence import NoteSequence
from diffusion_piano_b200.music.smf import parse_smf, write_smf


@pytest.mark.parametrize("factor", [0.5, 1.0, 2.0])
def test_temporal_stretch(factor):
    midi = music.load("CMajorScaleTwoHands")
    s = midi.stretch(factor)
    assert s.n_notes == midi.n_notes
    assert s.duration == midi.duration * factor


@pytest.mark.parametrize("factor", [-1, 0])
def test_temporal_stretch_raises(factor):
    with pytest.raises(ValueError):
        music.load("CMajorScaleTwoHands").stretch(factor)


@pytest.mark.parametrize("amount", [-2, -1, 0, 1, 2])
def test_transpose(amount):
    midi = music.load("CMajorScaleTwoHands")
    t = midi.transpose(amount)
    assert t.n_notes == midi.n_notes
    assert [n.pitch for n in t.seq.notes] == [n.pitch + amount for n in midi.seq.notes]


def test_trim_silence():
    trimmed = music.load("TwinkleTwinkleRousseau").trim_silence()
    assert trimmed.seq.notes[0].start_time == 0.0


def test_piano_note():
    number = midi_file.note_name_to_midi_number("C4")
    note = midi_file.PianoNote.create(number=number, velocity=100)
    assert (note.number, note.velocity, note.name) == (number, 100, "C4")
    for bad in (dict(number=-1, velocity=1), dict(number=128, velocity=1), dict(number=60, velocity=-1),
                dict(number=60, velocity=128)):
        with pytest.raises(ValueError):
            midi_file.PianoNote.create(**bad)
    assert midi_file.midi_number_to_key_number(midi_file.key_number_to_midi_number(10)) == 10
    assert midi_file.note_name_to_key_number(midi_file.key_number_to_note_name(39)) == 39


def _restrike_midi(dt=0.01):
    seq = NoteSequence()
    c6 = midi_file.note_name_to_midi_number("C6")
    seq.notes.add(start_time=1 * dt, end_time=2 * dt, velocity=80, pitch=c6, part=-1)
    seq.notes.add(start_time=2 * dt, end_time=5 * dt, velocity=80, pitch=c6, part=-1)
    seq.total_time = 5 * dt
    return midi_file.MidiFile(seq=seq)


def test_same_note_pressed_consecutively():  # midi_file_test.py:176-195
    traj = midi_file.NoteTrajectory.from_midi(_restrike_midi(), dt=0.01)
    assert len(traj) == 6
    assert traj.notes[0] == []
    assert [n.name for n in traj.notes[1]] == ["C6"]
    assert traj.notes[2] == []
    assert [n.name for n in traj.notes[3]] == ["C6"]
    assert [n.name for n in traj.notes[4]] == ["C6"]


def test_sustain():  # midi_file_test.py:197-211
    dt = 0.01
    seq = NoteSequence()
    c6 = midi_file.note_name_to_midi_number("C6")
    seq.notes.add(start_time=0, end_time=dt, velocity=80, pitch=c6, part=-1)
    seq.control_changes.add(time=0, control_number=64, control_value=64)
    seq.control_changes.add(time=3 * dt, control_number=64, control_value=0)
    seq.notes.add(start_time=5 * dt, end_time=6 * dt, velocity=80, pitch=c6, part=-1)
    seq.total_time = 6 * dt
    traj = midi_file.NoteTrajectory.from_midi(midi_file.MidiFile(seq=seq), dt=dt)
    assert len(traj) == 7
    assert all(traj.sustains[:3]) and not any(traj.sustains[3:6])


def test_bundled_songs_match_survey_numbers():
    tw = music.load("TwinkleTwinkleRousseau")
    assert tw.n_notes == 34 and tw.has_fingering() and len(tw.seq.control_changes) == 622
    assert abs(tw.duration - 22.546875) < 1e-9 and abs(tw.seq.notes[0].start_time - 14.6927083) < 1e-6
    assert len(midi_file.NoteTrajectory.from_midi(tw, 0.01)) == 2255
    traj = midi_file.NoteTrajectory.from_midi(tw.trim_silence(), 0.05)
    assert len(traj) == 158 and max(len(f) for f in traj.notes) == 4  # == len(twinkle_twinkle_actions.npy)
    gu = music.load("GurenNoYumiyaCut14s")
    assert gu.n_notes == 214 and gu.has_fingering()
    assert sorted({n.part for n in gu.seq.notes}) == list(range(1, 10))
    tg = midi_file.NoteTrajectory.from_midi(gu.trim_silence(), 0.05)
    assert len(tg) == 287 and max(len(f) for f in tg.notes) == 8
    cf = music.load("CrossingFieldCut10s")
    assert cf.n_notes == 18 and not cf.has_fingering()  # annotation file is in ticks: nothing matches
    assert len(midi_file.NoteTrajectory.from_midi(cf.trim_silence(), 0.05)) == 166


def test_smf_roundtrip_and_note_off_pairing():
    seq = NoteSequence()
    seq.tempos.add(qpm=120)
    seq.notes.add(pitch=60, velocity=70, start_time=0.0, end_time=0.5)
    seq.notes.add(pitch=64, velocity=71, start_time=0.25, end_time=1.0)
    seq.control_changes.add(time=0.5, control_number=64, control_value=127)
    back = parse_smf(write_smf(seq, ticks_per_quarter=480))
    assert [(n.pitch, n.velocity) for n in back.notes] == [(60, 70), (64, 71)]
    np.testing.assert_allclose([n.start_time for n in back.notes], [0.0, 0.25], atol=1e-9)
    np.testing.assert_allclose([n.end_time for n in back.notes], [0.5, 1.0], atol=1e-9)
    assert back.total_time == 1.0 and back.control_changes[0].control_value == 127


def test_song_tables_and_fingering_rule():
    traj = midi_file.NoteTrajectory.from_midi(music.load("CMajorChordProgressionTwoHands"), 0.05)
    tab = midi_file.SongTables.from_trajectory(traj)
    assert tab.goal.shape == (81, 89) and tab.pmax == 4
    assert tab.goal[0, :88].sum() == 4 and set(tab.pairs[0, :, 1]) == {5, 0, 2, 4}
    from diffusion_piano_b200.music.fingering import apply_annotation_rows, parse_pitch_to_midi_number
    assert parse_pitch_to_midi_number("C#1") == 25 and parse_pitch_to_midi_number("Bb3") == 58
    seq = NoteSequence()
    seq.notes.add(pitch=25, start_time=0.004, end_time=0.33)
    seq.notes.add(pitch=25, start_time=0.5, end_time=0.8)
    n = apply_annotation_rows(seq, [["0", "0.0", "0.33", "C#1", "50", "50", "1", "9"],
                                    ["1", "0.5", "0.8", "C#1", "50", "50", "1", "12"]])
    assert n == 1 and [x.part for x in seq.notes] == [9, 0]
