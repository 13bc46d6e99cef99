"""Task descriptions: the reference's task constructors, minus MuJoCo.

`SelfActuatedPiano` mirrors reference `robopianist/suite/tasks/self_actuated_piano.py:73-130`
and `PianoWithShadowHands` mirrors `piano_with_shadow_hands.py:50-128` + `base.py:96-135`
(same keyword names and defaults). A task object owns the compiled model and the song
tables; the physics and task logic themselves run in the CUDA step (`engine.py`).
"""

from __future__ import annotations

import enum
from collections import OrderedDict
from typing import List, Optional, Tuple

import numpy as np

from .. import _lib
from ..engine import EngineConfig
from ..music import midi_file
from ..music.midi_file import NoteTrajectory, SongTables
from ..scene import compile as scene_compile

_PHYSICS_TIMESTEP = 0.005  # reference suite/tasks/base.py:28
_CONTROL_TIMESTEP = 0.05  # :31
_ENERGY_PENALTY_COEF = 5e-3  # piano_with_shadow_hands.py:40
NUM_KEYS = 88


class RewardType(enum.Enum):
    NEGATIVE_XENT = "negative_xent"
    NEGATIVE_L2 = "negative_l2"


class _PianoTaskBase:
    """Timesteps + trajectory handling shared by both tasks (base.py:45-70)."""

    def __init__(self, midi, n_steps_lookahead, trim_silence, physics_timestep, control_timestep,
                 initial_buffer_time=0.0, augmentations=None):
        self._augmentations = list(augmentations) if augmentations is not None else None
        if trim_silence:
            midi = midi.trim_silence()
        self._midi = midi
        self._initial_midi = midi
        self._n_steps_lookahead = int(n_steps_lookahead)
        self.physics_timestep = float(physics_timestep)
        self.control_timestep = float(control_timestep)
        n = round(control_timestep / physics_timestep)
        if abs(n * physics_timestep - control_timestep) > 1e-9 * max(1.0, control_timestep) or n < 1:
            raise ValueError("control_timestep must be an integer multiple of physics_timestep")
        self.physics_steps_per_control_step = int(n)
        self._initial_buffer_time = initial_buffer_time
        self._reset_trajectory()

    def _reset_trajectory(self) -> None:
        traj = NoteTrajectory.from_midi(self._midi, self.control_timestep)
        traj.add_initial_buffer_time(self._initial_buffer_time)
        self._notes = traj.notes
        self._sustains = traj.sustains
        self.song = SongTables.from_trajectory(traj)

    def draw_song(self, random_state: np.random.RandomState):
        """`_maybe_change_midi` + `_reset_trajectory` of the reference (piano_with_shadow_hands.py:151-165) for ONE
        episode: the augmentations applied in order to the initial MIDI -> (midi, notes, sustains, SongTables).
        The batched environment draws a bank of these per reset and assigns one to every env."""
        midi = self._initial_midi
        for var in self._augmentations or ():
            midi = var(initial_value=midi, random_state=random_state)
        traj = NoteTrajectory.from_midi(midi, self.control_timestep)
        traj.add_initial_buffer_time(self._initial_buffer_time)
        return midi, traj.notes, traj.sustains, SongTables.from_trajectory(traj)

    @property
    def midi(self):
        return self._midi


class SelfActuatedPiano(_PianoTaskBase):
    def __init__(self, midi: midi_file.MidiFile, n_steps_lookahead: int = 0, trim_silence: bool = False,
                 reward_type: RewardType = RewardType.NEGATIVE_L2, augmentations=None,
                 change_color_on_activation: bool = False, physics_timestep: float = _PHYSICS_TIMESTEP,
                 control_timestep: float = _CONTROL_TIMESTEP) -> None:
        del change_color_on_activation  # visual only
        super().__init__(midi, n_steps_lookahead, trim_silence, physics_timestep, control_timestep,
                         augmentations=augmentations)
        self.reward_type = reward_type
        self.model = scene_compile.compile_scene(hands=False, piano_actuators=True,
                                                 physics_timestep=physics_timestep)
        self.reward_names = ("key_press_reward",)

    def engine_config(self, num_envs: int, device: int, canonical_actions: bool) -> EngineConfig:
        return EngineConfig(
            num_envs=num_envs, task=_lib.TASK_SELF_ACTUATED_PIANO,
            n_substeps=self.physics_steps_per_control_step, n_steps_lookahead=self._n_steps_lookahead,
            canonical_actions=canonical_actions,
            reward_type=(_lib.REWARD_NEGATIVE_L2 if self.reward_type == RewardType.NEGATIVE_L2
                         else _lib.REWARD_NEGATIVE_XENT), device=device)

    def observation_layout(self) -> "OrderedDict[str, Tuple[int, int]]":
        """name -> (offset, size) into the flat observation row the kernel writes."""
        g = (self._n_steps_lookahead + 1) * (NUM_KEYS + 1)
        # dict order follows composer: entity observables, then task observables
        return OrderedDict([("piano/activation", (g, NUM_KEYS)), ("piano/sustain_activation", (g + NUM_KEYS, 1)),
                            ("goal", (0, g))])


class PianoWithShadowHands(_PianoTaskBase):
    def __init__(self, midi: midi_file.MidiFile, n_steps_lookahead: int = 1,
                 n_seconds_lookahead: Optional[float] = None, trim_silence: bool = False,
                 wrong_press_termination: bool = False, initial_buffer_time: float = 0.0,
                 disable_fingering_reward: bool = False, disable_forearm_reward: bool = False,
                 disable_colorization: bool = False, disable_hand_collisions: bool = False,
                 augmentations=None, energy_penalty_coef: float = _ENERGY_PENALTY_COEF,
                 randomize_hand_positions: bool = False, gravity_compensation: bool = False,
                 change_color_on_activation: bool = False, primitive_fingertip_collisions: bool = False,
                 reduced_action_space: bool = False, attachment_yaw: float = 0.0,
                 forearm_dofs=("forearm_tx", "forearm_ty"), physics_timestep: float = _PHYSICS_TIMESTEP,
                 control_timestep: float = _CONTROL_TIMESTEP) -> None:
        del disable_colorization, change_color_on_activation  # visual only
        # primitive_fingertip_collisions: the mesh colliders are unavailable, capsules are always used
        del primitive_fingertip_collisions
        # attachment_yaw: any angle; forearm_dofs: subsets of the default pair (others raise in compile_scene)
        self._reduced_action_space = bool(reduced_action_space)
        self._randomize_hand_positions = bool(randomize_hand_positions)
        if n_seconds_lookahead is not None:
            n_steps_lookahead = int(np.ceil(n_seconds_lookahead / control_timestep))
        super().__init__(midi, n_steps_lookahead, trim_silence, physics_timestep, control_timestep,
                         initial_buffer_time=initial_buffer_time, augmentations=augmentations)
        self._disable_fingering_reward = disable_fingering_reward or not self._midi.has_fingering()
        self._disable_forearm_reward = disable_forearm_reward
        self._wrong_press_termination = wrong_press_termination
        self._disable_hand_collisions = disable_hand_collisions
        self._energy_penalty_coef = energy_penalty_coef
        self.model = scene_compile.compile_scene(
            hands=True, gravity_compensation=gravity_compensation,
            disable_hand_collisions=disable_hand_collisions, physics_timestep=physics_timestep,
            reduced_action_space=self._reduced_action_space, attachment_yaw=float(attachment_yaw),
            forearm_dofs=tuple(forearm_dofs))
        names = ["key_press_reward", "sustain_reward", "energy_reward"]
        names.append("ot_fingering_reward" if self._disable_fingering_reward else "fingering_reward")
        if not disable_forearm_reward:
            names.append("forearm_reward")
        self.reward_names = tuple(names)

    def engine_config(self, num_envs: int, device: int, canonical_actions: bool) -> EngineConfig:
        return EngineConfig(
            num_envs=num_envs, task=_lib.TASK_PIANO_WITH_SHADOW_HANDS,
            n_substeps=self.physics_steps_per_control_step, n_steps_lookahead=self._n_steps_lookahead,
            canonical_actions=canonical_actions, has_fingering=not self._disable_fingering_reward,
            disable_forearm_reward=self._disable_forearm_reward,
            wrong_press_termination=self._wrong_press_termination,
            energy_penalty_coef=self._energy_penalty_coef, device=device)

    def reduced_spec(self):
        """reduced_action_space (shadow_hand.py:160-181) and forearm_dofs subsets (shadow_hand.py:272-311): (kept action columns of the engine's 45, kept observation
        columns of the engine's row, layout of the compacted row) - the removed joints' three actuators and three
        joint positions per hand leave both specs. None for the full hand."""
        if not self.model.locked_dofs:
            return None
        full = self.observation_layout(full=True)
        act_keep = [i for i in range(45) if i not in self.model.dead_actuators]
        drop = set()
        for h, key in enumerate(("rh_shadow_hand/joints_pos", "lh_shadow_hand/joints_pos")):
            off, _ = full[key]
            for d in self.model.locked_dofs:
                lane = d - NUM_KEYS - 26 * h
                if 0 <= lane < 26:
                    drop.add(off + (lane - 2 if lane >= 2 else 24 + lane))  # hand.joints order: 24 joints, then forearm
        width = max(o + n for o, n in full.values())
        obs_keep = [c for c in range(width) if c not in drop]
        new_of = {c: i for i, c in enumerate(obs_keep)}
        layout = OrderedDict()
        for k, (o, n) in full.items():
            cols = [new_of[c] for c in range(o, o + n) if c in new_of]
            layout[k] = (cols[0], len(cols))
        return act_keep, obs_keep, layout

    def observation_layout(self, full: bool = False) -> "OrderedDict[str, Tuple[int, int]]":
        if self.model.locked_dofs and not full:
            return self.reduced_spec()[2]
        g = (self._n_steps_lookahead + 1) * (NUM_KEYS + 1)
        f = 0 if self._disable_fingering_reward else 10
        o = g + f
        layout = OrderedDict([("piano/state", (o, NUM_KEYS)), ("piano/sustain_state", (o + NUM_KEYS, 1)),
                              ("rh_shadow_hand/joints_pos", (o + NUM_KEYS + 1, 26)),
                              ("lh_shadow_hand/joints_pos", (o + NUM_KEYS + 27, 26)), ("goal", (0, g))])
        if f:
            layout["fingering"] = (g, f)
        return layout
