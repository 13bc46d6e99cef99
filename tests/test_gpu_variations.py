"""MIDI augmentations on the device (reference piano_with_shadow_hands.py:151-165 applying
suite/variations.py at every episode start): a bank of augmented songs resident in HBM, one per env. Every env is
checked against a CPU oracle built for ITS song: goal / fingering observations bit-equal, episode length, reward."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _augmented_task():
    from diffusion_piano_b200 import music, suite
    from diffusion_piano_b200.suite import variations as V
    return suite.PianoWithShadowHands(
        music.load("CMajorChordProgressionTwoHands"), n_steps_lookahead=2, control_timestep=0.05,
        augmentations=[V.MidiTemporalStretch(prob=0.8, stretch_range=0.3), V.MidiPitchShift(prob=0.8, shift_range=4),
                       V.MidiOctaveShift(prob=0.3, octave_range=1)])


def test_every_env_plays_its_own_augmented_song():
    from diffusion_piano_b200 import suite
    from oracle.task_oracle import PianoWithShadowHandsOracle
    task = _augmented_task()
    n = 5
    env = suite.BatchedEnvironment(task, n, canonical_actions=False, auto_reset=False, random_state=11,
                                   augmentation_bank_size=n)
    for episode in range(2):
        obs = env.reset()
        assert len(env.songs) == n and list(env.song_ids) == list(range(n))
        frames = [env.songs[i][3].n_frames for i in range(n)]
        assert len(set(frames)) > 1 or len({tuple(env.songs[i][3].goal[0]) for i in range(n)}) > 1
        oras = []
        for i in range(n):  # the same task on env i's song, through the CPU oracle
            t_i = suite.PianoWithShadowHands(env.songs[i][0], n_steps_lookahead=2, control_timestep=0.05)
            assert len(t_i._notes) == frames[i]
            oras.append(PianoWithShadowHandsOracle(t_i))
            o0 = oras[-1].reset()
            np.testing.assert_array_equal(obs["goal"][i], o0["goal"])
            np.testing.assert_array_equal(obs["fingering"][i], o0["fingering"])
        a = np.zeros(45)
        a[:44] = 0.5 * (oras[0].act_lo + oras[0].act_hi)
        alive = [True] * n
        for step in range(max(frames)):
            obs, rew, done = env.step(np.stack([a] * n))
            for i in range(n):
                if not alive[i]:
                    continue
                obs_o, r_o, d_o = oras[i].step(np.asarray(a, np.float32).astype(np.float64))
                assert bool(done[i]) == d_o == (step == frames[i] - 1), (episode, i, step)
                np.testing.assert_array_equal(obs["goal"][i], obs_o["goal"], err_msg=f"env {i} step {step}")
                np.testing.assert_array_equal(obs["fingering"][i], obs_o["fingering"])
                np.testing.assert_allclose(rew[i], r_o, atol=5e-3)
                alive[i] = not d_o
        assert not any(alive)
    env.close()


def test_auto_reset_envs_take_a_bank_entry_and_the_dm_env_task_view_follows():
    import torch
    from diffusion_piano_b200 import suite
    task = _augmented_task()
    n = 64
    env = suite.BatchedEnvironment(task, n, canonical_actions=True, auto_reset=True, random_state=2,
                                   augmentation_bank_size=8)
    env.reset_device()
    assert len(env.songs) == 8 and env.song_ids.min() >= 0 and env.song_ids.max() < 8
    frames = np.array([env.songs[k][3].n_frames for k in env.song_ids])
    a = torch.zeros((n, 45), device="cuda")
    first_done = np.full(n, -1)
    for step in range(int(frames.max()) + 1):
        _, _, done, _, _ = env.step_device(a)
        d = done.cpu().numpy().astype(bool)
        first_done[(first_done < 0) & d] = step
    np.testing.assert_array_equal(first_done, frames - 1)  # every env ran exactly its own song's length
    env.close()
    # dm_env view: the task object carries the episode's song (MidiEvaluationWrapper reads task._notes)
    e1 = suite.Environment(_augmented_task(), random_state=5)
    lengths = set()
    for _ in range(4):
        ts = e1.reset()
        steps = 0
        while not ts.last():
            ts = e1.step(np.zeros(45))
            steps += 1
        assert steps == len(e1.task._notes) == e1.task.song.n_frames
        lengths.add(steps)
    assert len(lengths) > 1
    e1.close()
