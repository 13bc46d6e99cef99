"""MidiModule (reference models/piano/midi_module.py:47-98): the host replay rules on CPU, and on the GPU the
per-substep activation words of the step kernel against the oracle's key state after every substep."""
import numpy as np
import pytest


def test_event_rules_cpu():
    from diffusion_piano_b200.music.midi_module import MidiModule, NoteOff, NoteOn, SustainOff, SustainOn, \
        unpack_activation_words

    words = np.zeros((4, 4), np.uint32)
    words[0] = (1 << 3, 0, 0, 0)                 # key 3 on
    words[1] = (1 << 3, 1 << 1, 1 << 23, 1)      # + keys 33, 87, pedal
    words[2] = (0, 1 << 1, 1 << 23, 1)           # key 3 off
    words[3] = (0, 0, 0, 0)                      # everything off
    act, sus = unpack_activation_words(words)
    assert act.shape == (4, 88) and list(np.flatnonzero(act[1])) == [3, 33, 87] and list(sus) == [False, True, True, False]
    m = MidiModule()
    got = []
    m.register_synth_note_on_callback(lambda n, v: got.append(("on", n, v)))
    m.register_synth_note_off_callback(lambda n: got.append(("off", n)))
    m.register_synth_sustain_on_callback(lambda: got.append("sus_on"))
    m.register_synth_sustain_off_callback(lambda: got.append("sus_off"))
    m.after_step_words(words, first_substep_index=10, timestep=0.005)
    ev = m.get_all_midi_messages()
    assert ev[0] == NoteOn(note=24, velocity=127, time=pytest.approx(0.055))
    assert ev[1:4] == [NoteOn(54, 127, pytest.approx(0.060)), NoteOn(108, 127, pytest.approx(0.060)),
                       SustainOn(pytest.approx(0.060))]
    assert ev[4] == NoteOff(24, pytest.approx(0.065))
    assert ev[5:] == [NoteOff(54, pytest.approx(0.070)), NoteOff(108, pytest.approx(0.070)), SustainOff(pytest.approx(0.070))]
    assert m.get_latest_midi_messages() == ev[5:]
    assert got == [("on", 24, 127), ("on", 54, 127), ("on", 108, 127), "sus_on", ("off", 24), ("off", 54),
                   ("off", 108), "sus_off"]
    m.initialize_episode()
    assert m.get_all_midi_messages() == []


@pytest.mark.gpu
def test_substep_activations_match_oracle():
    """Keys pushed down by an applied torque (as piano_with_shadow_hands_test.py:228-242 does): the words recorded
    after every substep equal the oracle's `Piano._update_key_state` after every `mj_step`, and the events follow."""
    import torch

    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.music.midi_module import MidiModule, unpack_activation_words
    from oracle.task_oracle import PianoWithShadowHandsOracle

    task = suite.PianoWithShadowHands(music.load("CMajorScaleTwoHands"), n_steps_lookahead=1, control_timestep=0.05)
    env = suite.BatchedEnvironment(task, 2, auto_reset=False)
    mods = env.enable_midi_events((0, 1))
    ora = PianoWithShadowHandsOracle(task)
    env.reset()
    ora.reset()
    ref = MidiModule()
    nsub, h = task.physics_steps_per_control_step, task.physics_timestep
    push = np.zeros(140)
    rs = np.random.RandomState(0)
    sub_index = 0
    for step in range(6):
        push[:88] = np.where(rs.rand(88) < 0.3, 0.6, 0.0) if step % 2 == 0 else 0.0  # press a third of the keys, release
        f = torch.as_tensor(np.stack([push, push]).astype(np.float32), device="cuda")
        env.engine.set_qfrc_applied(f)
        ora.physics.qfrc_applied[:] = push
        a = np.zeros(45)
        a[44] = 1.0 if step in (2, 3) else 0.0  # pedal
        # the oracle, substep by substep
        p = ora.physics
        p.ctrl[:] = a[:-1]
        ora.sustain_state = a[-1]
        for _ in range(nsub):
            p.step()
            ora._key_state()
            sub_index += 1
            ref.after_substep(sub_index * h, ora.activation, ora.sustain_activation)
        p.forward_position(True)
        env.step(np.stack([a, a]))
        words = env._midi_words.cpu().numpy().view(np.uint32)
        act, sus = unpack_activation_words(words[0])
        np.testing.assert_array_equal(words[0], words[1])
        assert sus.all() == (step in (2, 3))
    ev_ref, ev = ref.get_all_midi_messages(), mods[0].get_all_midi_messages()
    assert len(ev_ref) > 20 and any(type(e).__name__ == "NoteOff" for e in ev_ref)
    assert [(type(e).__name__, getattr(e, "note", -1)) for e in ev] == \
        [(type(e).__name__, getattr(e, "note", -1)) for e in ev_ref]
    np.testing.assert_allclose([e.time for e in ev], [e.time for e in ev_ref], atol=1e-9)
    assert mods[1].get_all_midi_messages() == ev
    env.close()
