"""Environment façades over the CUDA step.

* `BatchedEnvironment`  - N envs on one GPU; device-tensor API (`step_device`) for training
  loops that stay on the GPU, and the reference's numpy dict API (`reset`/`step`).
* `VectorizedPianoEnv`  - drop-in for reference `parallelized_base_v2.py:21-67` (boundary B1).
* `Environment`         - single-env dm_env API of `composer_utils.Environment(task=...)` /
  `robopianist.suite.load` (boundary B2): reset()/step() -> TimeStep, FIRST has reward None.
"""

from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Optional

import numpy as np

from ..engine import BatchedPianoEngine
from . import specs


class BatchedEnvironment:
    def __init__(self, task, num_envs: int, device: int = 0, canonical_actions: bool = False,
                 auto_reset: bool = True, max_contacts: int = 0, random_state=None, schedule: str = "auto",
                 augmentation_bank_size: int = 64, legacy_step: bool = True):
        import torch  # device memory + streams only

        self._torch = torch
        self.task = task
        self.num_envs = int(num_envs)
        cfg = task.engine_config(num_envs, device, canonical_actions)
        cfg.auto_reset = auto_reset
        self._auto_reset = bool(auto_reset)
        cfg.max_contacts = max_contacts
        # dm_control Physics.legacy_step (composer.Environment(legacy_step=), suite.load): False = plain mj_step, the task
        # then reads positions, contacts and actuator velocities of the last substep's START
        cfg.legacy_step = self._legacy_step = bool(legacy_step)
        self.engine = BatchedPianoEngine(task.model, task.song, cfg)
        if schedule != "auto":  # launch shape of the two-hands step (engine.set_schedule); never changes results
            self.engine.set_schedule(schedule)
        self.device = torch.device("cuda", device)
        self.layout = task.observation_layout()
        self.obs_dim, self.action_dim = self.engine.obs_dim, self.engine.action_dim
        # reduced_action_space: the engine keeps its 45 actions / full observation row; the removed joints' columns
        # are dropped (observations) and zero-filled (actions: their actuators are dead) at this boundary
        red = task.reduced_spec() if hasattr(task, "reduced_spec") else None
        self._reduced = red is not None
        self._full_obs_dim, self._full_action_dim = self.obs_dim, self.action_dim
        self._pos_map = None
        if self._reduced:
            self._act_keep, self._obs_keep = np.asarray(red[0]), np.asarray(red[1])
            self.obs_dim, self.action_dim = len(self._obs_keep), len(self._act_keep)
            if hasattr(task, "position_map"):  # one-hand task: `<hand>/position`, affine in two joints_pos columns
                cols, W, c = task.position_map()
                self._pos_map = (np.asarray(cols, np.int64), np.asarray(W, np.float64), np.asarray(c, np.float64))
                self.obs_dim += 3
        n = self.num_envs
        self._obs = torch.zeros((n, self._full_obs_dim), dtype=torch.float32, device=self.device)
        if self._reduced:
            self._act_keep_t = torch.as_tensor(self._act_keep, device=self.device)
            self._obs_keep_t = torch.as_tensor(self._obs_keep, device=self.device)
            self._obs_red = torch.zeros((n, self.obs_dim), dtype=torch.float32, device=self.device)
            if self._pos_map is not None:
                self._obs_sel = torch.zeros((n, len(self._obs_keep)), dtype=torch.float32, device=self.device)
                self._pos_cols_t = torch.as_tensor(self._pos_map[0], device=self.device)
                self._pos_W_t = torch.as_tensor(self._pos_map[1], dtype=torch.float32, device=self.device)
                self._pos_c_t = torch.as_tensor(self._pos_map[2], dtype=torch.float32, device=self.device)
                if not self._legacy_step:  # the root body's xpos is one substep old: q - h v of the forearm slides
                    self._pos_dofs_t = torch.as_tensor(np.asarray(task.position_dofs(), np.int64), device=self.device)
                    self._qvel = torch.zeros((n, self.engine.nv), dtype=torch.float32, device=self.device)
            self._act_full = torch.zeros((n, self._full_action_dim), dtype=torch.float32, device=self.device)
        self._reward = torch.zeros(n, dtype=torch.float32, device=self.device)
        self._terms = torch.zeros((n, 5), dtype=torch.float32, device=self.device)
        self._discount = torch.ones(n, dtype=torch.float32, device=self.device)
        self._done = torch.zeros(n, dtype=torch.uint8, device=self.device)
        self._canonical = canonical_actions
        # host side of the numpy API: page-locked, so pp_step_host copies straight into these arrays
        def pinned(shape, dtype, fill=0):
            t = torch.full(shape, fill, dtype=dtype).pin_memory()
            self._pinned.append(t)  # keeps the allocation alive under the numpy view
            return t.numpy()

        self._pinned = []
        self._h_obs = pinned((n, self._full_obs_dim), torch.float32)
        self._h_reward = pinned((n,), torch.float32)
        self._h_terms = pinned((n, 5), torch.float32)
        self._h_discount = pinned((n,), torch.float32, 1)
        self._h_done = pinned((n,), torch.uint8)
        self._h_actions = pinned((n, self._full_action_dim), torch.float32)
        # randomize_hand_positions (piano_with_shadow_hands.py:491-499): a uniform(-0.05, 0.05) y shift of
        # both hands, drawn per env and per episode from the environment's RandomState
        self._randomize = bool(getattr(task, "_randomize_hand_positions", False))
        if self._randomize:
            if isinstance(random_state, np.random.RandomState):
                self.random_state = random_state
            else:
                self.random_state = np.random.RandomState(random_state)
            self._h_offsets = pinned((n,), torch.float32)
            self._d_offsets = torch.zeros(n, dtype=torch.float32, device=self.device)

        # MIDI augmentations (piano_with_shadow_hands.py:151-165, suite/variations.py): a bank of K augmented songs
        # is drawn and uploaded at every explicit reset; every env takes one of them (an env that resets itself
        # between two explicit resets draws again from the same bank)
        self._augment = bool(getattr(task, "_augmentations", None))
        if self._augment:
            if not hasattr(self, "random_state"):
                self.random_state = (random_state if isinstance(random_state, np.random.RandomState)
                                     else np.random.RandomState(random_state))
            self._bank_size = max(1, min(int(augmentation_bank_size), n))
            self._d_song_ids = torch.zeros(n, dtype=torch.int32, device=self.device)
            self.songs = []  # the bank in force: [(midi, notes, sustains, SongTables)]
            self.song_ids = np.zeros(n, np.int32)  # bank entry of every env's current episode

    _POSITION_OFFSET = 0.05

    def _draw_songs(self, new_bank: bool):
        """new_bank: draw K fresh augmented songs (explicit reset). Always: hand every env the bank entry it takes
        at its next reset - env i plays entry i of a fresh bank when the bank is as large as the batch (the
        reference's one independent draw per env and episode), a uniformly drawn entry otherwise."""
        if not self._augment:
            return
        stream = self._torch.cuda.current_stream(self.device)
        if new_bank:
            self.songs = [self.task.draw_song(self.random_state) for _ in range(self._bank_size)]
            stream.synchronize()
            self.engine.set_song_bank([s[3] for s in self.songs])
        if new_bank and self._bank_size == self.num_envs:
            ids = np.arange(self.num_envs, dtype=np.int32)
        else:
            ids = self.random_state.randint(0, self._bank_size, self.num_envs).astype(np.int32)
        self._pending_song_ids = ids
        self._d_song_ids.copy_(self._torch.from_numpy(ids))
        self.engine.set_song_ids(self._d_song_ids, stream.cuda_stream)
        stream.synchronize()
        if new_bank:
            self.song_ids = ids.copy()

    def _draw_hand_offsets(self):
        """Offsets the envs take at their NEXT reset; redrawn before every call that can reset an env."""
        if not self._randomize:
            return
        self._h_offsets[...] = self.random_state.uniform(-self._POSITION_OFFSET, self._POSITION_OFFSET,
                                                         self.num_envs)
        self._d_offsets.copy_(self._torch.from_numpy(self._h_offsets), non_blocking=True)
        stream = self._torch.cuda.current_stream(self.device).cuda_stream
        self.engine.set_hand_offsets(self._d_offsets, stream)
        self._torch.cuda.current_stream(self.device).synchronize()

    # ---- specs
    def action_spec(self) -> specs.BoundedArray:
        if self._canonical:
            return specs.BoundedArray((self.action_dim,), np.float64, -1.0, 1.0, name="action")
        lo, hi = self.engine.action_minimum, self.engine.action_maximum
        if self._reduced:
            lo, hi = lo[self._act_keep], hi[self._act_keep]
        return specs.BoundedArray((self.action_dim,), np.float64, lo, hi, name="action")

    def observation_spec(self) -> "OrderedDict[str, specs.Array]":
        return OrderedDict((k, specs.Array((size,), np.float64, name=k)) for k, (_, size) in self.layout.items())

    # ---- device API: everything stays in HBM; returned tensors are reused between calls
    def reset_device(self, mask=None):
        self._draw_hand_offsets()
        self._draw_songs(new_bank=mask is None)
        stream = self._torch.cuda.current_stream(self.device).cuda_stream
        self.engine.reset(obs=self._obs, mask=mask, stream=stream)
        return self._device_obs()

    def _device_obs(self):
        if not self._reduced:
            return self._obs
        if self._pos_map is None:
            self._torch.index_select(self._obs, 1, self._obs_keep_t, out=self._obs_red)
            return self._obs_red
        k = len(self._obs_keep)
        self._torch.index_select(self._obs, 1, self._obs_keep_t, out=self._obs_sel)
        self._obs_red[:, :k].copy_(self._obs_sel)
        fore = self._obs_sel.index_select(1, self._pos_cols_t)
        if not self._legacy_step:
            stream = self._torch.cuda.current_stream(self.device).cuda_stream
            self.engine.get_state(qvel=self._qvel, stream=stream)
            fore = fore - self.task.physics_timestep * self._qvel.index_select(1, self._pos_dofs_t)
        self._obs_red[:, k:] = self._pos_c_t + fore @ self._pos_W_t
        return self._obs_red

    def step_device(self, actions):
        """actions: float32 CUDA tensor (N, action_dim). Returns (obs, reward, done, terms, discount)."""
        if self._auto_reset:
            self._draw_hand_offsets()
        stream = self._torch.cuda.current_stream(self.device).cuda_stream
        if self._reduced:
            self._act_full.index_copy_(1, self._act_keep_t, actions)
            actions = self._act_full
        self.engine.step(actions, self._obs, self._reward, self._done, self._terms, self._discount, stream)
        return self._device_obs(), self._reward, self._done, self._terms, self._discount

    # ---- numpy API (host buffers; copies inside pp_step_host)
    def _split(self, flat: np.ndarray) -> Dict[str, np.ndarray]:
        if self._reduced:
            flat = flat[:, self._obs_keep]
            if self._pos_map is not None:
                cols, W, c = self._pos_map
                fore = flat[:, cols].astype(np.float64)
                if not self._legacy_step:
                    self.engine.get_state(qvel=self._qvel)
                    self._torch.cuda.synchronize(self.device)
                    fore = fore - self.task.physics_timestep * self._qvel[:, self._pos_dofs_t].double().cpu().numpy()
                flat = np.concatenate([flat, (c + fore @ W).astype(flat.dtype)], 1)
        return OrderedDict((k, flat[:, o:o + s].astype(np.float64)) for k, (o, s) in self.layout.items())

    def reset(self) -> Dict[str, np.ndarray]:
        self._draw_hand_offsets()
        self._draw_songs(new_bank=True)
        self.engine.reset_host(obs=self._h_obs)
        for i, m in getattr(self, "_midi_modules", {}).items():
            m.initialize_episode()
            self._midi_substeps[i] = 0
        self._h_done[...] = 0
        return self._split(self._h_obs)

    # ---- MIDI events (reference Piano.midi_module, models/piano/midi_module.py): recorded per substep on the device
    def enable_midi_events(self, envs=(0,)):
        """Starts recording the key / pedal activations after every physics substep (pp_set_substep_activations)
        and returns {env index: MidiModule}; the modules are fed by the numpy `step()` of this class."""
        from ..music.midi_module import MidiModule

        torch = self._torch
        nsub = self.task.physics_steps_per_control_step
        if not hasattr(self, "_midi_words"):
            self._midi_words = torch.zeros((self.num_envs, nsub, 4), dtype=torch.int32, device=self.device)
            self.engine.set_substep_activations(self._midi_words)
            self._midi_modules, self._midi_substeps = {}, {}
        for i in envs:
            if i not in self._midi_modules:
                self._midi_modules[int(i)] = MidiModule()
                self._midi_substeps[int(i)] = 0
        return self._midi_modules

    def _feed_midi(self, was_done: np.ndarray) -> None:
        idx = sorted(self._midi_modules)
        words = self._midi_words[idx].cpu().numpy().view(np.uint32)
        nsub = words.shape[1]
        for row, i in enumerate(idx):
            m = self._midi_modules[i]
            if self._auto_reset and was_done[i]:  # this call reset the env: no substep ran (initialize_episode)
                m.initialize_episode()
                self._midi_substeps[i] = 0
                continue
            m.after_step_words(words[row], self._midi_substeps[i], self.task.physics_timestep)
            self._midi_substeps[i] += nsub

    def step(self, actions):
        a = np.asarray(actions)
        if a.shape != (self.num_envs, self.action_dim):
            raise ValueError(f"actions must have shape {(self.num_envs, self.action_dim)}, got {a.shape}")
        if self._reduced:
            self._h_actions[:, self._act_keep] = a
        else:
            self._h_actions[...] = a  # float64 -> float32 into the page-locked buffer
        if self._auto_reset:
            self._draw_hand_offsets()
        was_done = self._h_done.astype(bool) if getattr(self, "_midi_modules", None) else None
        self.engine.step_host(self._h_actions, self._h_obs, self._h_reward, self._h_done, self._h_terms,
                              self._h_discount)
        if was_done is not None:
            self._feed_midi(was_done)
        return self._split(self._h_obs), self._h_reward.astype(np.float64), self._h_done.astype(bool)

    @property
    def reward_terms(self) -> Dict[str, np.ndarray]:
        """Per-term rewards of the last numpy step (CompositeReward.reward_terms, per env)."""
        cols = {"key_press_reward": 0, "sustain_reward": 1, "energy_reward": 2, "fingering_reward": 3,
                "ot_fingering_reward": 3, "forearm_reward": 4}
        return OrderedDict((n, self._h_terms[:, cols[n]].astype(np.float64)) for n in self.task.reward_names)

    # ---- musical metrics (wrappers/evaluation.py:67-177), accumulated by the step on the device
    def episode_metrics(self):
        """(metrics [N,6] float64, episodes [N] int): means over the steps of each env's last finished
        episode {precision, recall, f1, sustain_precision, sustain_recall, sustain_f1} and the number of
        episodes the env has finished."""
        torch = self._torch
        m = torch.zeros((self.num_envs, 6), dtype=torch.float32, device=self.device)
        c = torch.zeros(self.num_envs, dtype=torch.int32, device=self.device)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        self.engine.get_metrics(m, c, stream)
        return m.cpu().numpy().astype(np.float64), c.cpu().numpy()

    def get_musical_metrics(self) -> Dict[str, float]:
        """Batched counterpart of MidiEvaluationWrapper.get_musical_metrics: mean over the envs' last
        finished episodes."""
        m, c = self.episode_metrics()
        if not (c > 0).any():
            raise ValueError("No episode metrics available yet.")
        names = ("precision", "recall", "f1", "sustain_precision", "sustain_recall", "sustain_f1")
        return {n: float(m[c > 0, j].mean()) for j, n in enumerate(names)}

    def close(self):
        self.engine.close()


class _SpecView:
    """What `vec_env.envs[0]` must answer (parallelized_base_v2.py:82-86)."""

    def __init__(self, batched: BatchedEnvironment):
        self._b = batched

    def observation_spec(self):
        return self._b.observation_spec()

    def action_spec(self):
        return self._b.action_spec()


class VectorizedPianoEnv:
    """Same constructor and methods as reference parallelized_base_v2.py:21-67; the N envs are
    one batched CUDA step instead of a Python loop. Task kwargs are the reference's (:28-38);
    actions are canonical [-1, 1] (CanonicalSpecWrapper, :44)."""

    def __init__(self, num_envs, midi_sequence, device: int = 0, **task_overrides):
        from .tasks import PianoWithShadowHands

        kwargs = dict(n_steps_lookahead=1, trim_silence=True, wrong_press_termination=False,
                      initial_buffer_time=0.0, disable_fingering_reward=False, disable_forearm_reward=False,
                      disable_colorization=False, disable_hand_collisions=False)
        kwargs.update(task_overrides)
        self.num_envs = num_envs
        self.task = PianoWithShadowHands(midi=midi_sequence, **kwargs)
        self.batched = BatchedEnvironment(self.task, num_envs, device=device, canonical_actions=True,
                                          auto_reset=True)
        self.envs = [_SpecView(self.batched)] * num_envs

    def reset(self):
        return self.batched.reset()

    def step(self, actions):
        return self.batched.step(actions)


class Environment:
    """dm_env-style single environment (N = 1 batch)."""

    def __init__(self, task, device: int = 0, canonical_actions: bool = False, random_state=None,
                 strip_singleton_obs_buffer_dim: bool = True, recompile_physics: bool = False,
                 legacy_step: bool = True):
        del strip_singleton_obs_buffer_dim, recompile_physics
        self._b = BatchedEnvironment(task, 1, device=device, canonical_actions=canonical_actions,
                                     auto_reset=False, random_state=random_state, legacy_step=legacy_step)
        self.task = task
        self._reset_next_step = True

    def action_spec(self):
        return self._b.action_spec()

    def observation_spec(self):
        return self._b.observation_spec()

    def _obs(self, d):
        return OrderedDict((k, v[0]) for k, v in d.items())

    def reset(self) -> specs.TimeStep:
        self._reset_next_step = False
        self._t = 0
        obs = self._b.reset()
        if self._b._augment:  # the task object shows the episode's own song, as `_maybe_change_midi` leaves it
            midi, notes, sustains, tables = self._b.songs[int(self._b.song_ids[0])]
            self.task._midi, self.task._notes, self.task._sustains, self.task.song = midi, notes, sustains, tables
        return specs.TimeStep(specs.StepType.FIRST, None, None, self._obs(obs))

    def _cache_current(self) -> None:
        """What after_step caches for the reward of the frame just played (piano_with_shadow_hands.py:192-199):
        `task._goal_current`, `_rh_keys_current`, `_lh_keys_current`. The device holds the frame counter; these
        host mirrors are for callers that read the task object (the reference's tests, MidiEvaluationWrapper)."""
        song = getattr(self.task, "song", None)
        if song is None or self._t >= song.goal.shape[0]:
            return
        self.task._goal_current = np.asarray(song.goal[self._t], np.float64)
        pairs = getattr(song, "pairs", None)
        if pairs is not None:
            cur = [(int(k), int(f)) for k, f in np.asarray(pairs[self._t]) if k >= 0]
            self.task._rh_keys_current = [(k, f) for k, f in cur if f < 5]
            self.task._lh_keys_current = [(k, f - 5) for k, f in cur if f >= 5]

    def step(self, action) -> specs.TimeStep:
        if self._reset_next_step:
            return self.reset()
        obs, reward, done = self._b.step(np.asarray(action, np.float32)[None])
        self._cache_current()
        self._t += 1
        discount = float(self._b._h_discount[0])
        if done[0]:
            self._reset_next_step = True
            return specs.TimeStep(specs.StepType.LAST, float(reward[0]), discount, self._obs(obs))
        return specs.TimeStep(specs.StepType.MID, float(reward[0]), 1.0, self._obs(obs))

    @property
    def reward_terms(self):
        return OrderedDict((k, float(v[0])) for k, v in self._b.reward_terms.items())

    @property
    def batched(self) -> BatchedEnvironment:
        return self._b

    @property
    def midi_module(self):
        """`environment.task.piano.midi_module` of the reference (viewer/runtime.py:135): note / pedal events of
        this env, recorded from the first access on."""
        return self._b.enable_midi_events((0,))[0]

    def close(self):
        self._b.close()
