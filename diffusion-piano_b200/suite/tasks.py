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
            canonical_actions=canonical_actions, has_fingering=self._fingering_mode(),
            disable_forearm_reward=self._disable_forearm_reward,
            wrong_press_termination=self._wrong_press_termination,
            energy_penalty_coef=self._energy_penalty_coef, device=device)

    def _fingering_mode(self) -> int:
        """pp_config.has_fingering: 1 fingering reward + observation, 0 optimal-transport fingering reward."""
        return 0 if self._disable_fingering_reward else 1

    def _extra_dropped_columns(self, full) -> set:
        """Columns of the engine's observation row, besides the locked joints', that the task does not expose."""
        return set()

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
        drop |= self._extra_dropped_columns(full)
        width = max(o + n for o, n in full.values())
        obs_keep = [c for c in range(width) if c not in drop]
        new_of = {c: i for i, c in enumerate(obs_keep)}
        layout = OrderedDict()
        for k, (o, n) in full.items():
            cols = [new_of[c] for c in range(o, o + n) if c in new_of]
            if cols:  # an observable whose every column is gone (the absent hand's joints) is not in the spec
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


class HandSide(enum.Enum):
    """Which hand the one-hand task has (reference robopianist/models/hands/base.py:26-30)."""

    LEFT = enum.auto()
    RIGHT = enum.auto()


class PianoWithOneShadowHand(PianoWithShadowHands):
    """reference `robopianist/suite/tasks/piano_with_one_shadow_hand.py:40-388`: one hand plays the whole MIDI.

    Same keyword names and defaults as the reference (its `**kwargs` are `PianoTask`'s, base.py:96-107). Rewards:
    key_press, sustain, energy of THIS hand, and fingering over this hand's notes unless disabled (no
    optimal-transport variant, no forearm term: :110-118). Observations: goal, fingering (5), piano/state,
    piano/sustain_state, `<hand>/joints_pos`, `<hand>/position` (:320-361).
    The step kernel always carries two hands: the absent one is locked, parked and dead (scene/compile.py,
    `absent_hand`), which the fp64 checker shows equal to the reference's one-hand model to 1e-13; its joints, its
    actuators and its half of the fingering flags leave the specs at the facade."""

    def __init__(self, midi: midi_file.MidiFile, hand_side: HandSide, n_steps_lookahead: int = 1,
                 n_seconds_lookahead: Optional[float] = None, trim_silence: bool = False,
                 wrong_press_termination: bool = False, initial_buffer_time: float = 0.0,
                 disable_fingering_reward: bool = False, disable_colorization: bool = False, augmentations=None,
                 gravity_compensation: bool = False, change_color_on_activation: bool = False,
                 primitive_fingertip_collisions: bool = False, reduced_action_space: bool = False,
                 attachment_yaw: float = 0.0, forearm_dofs=("forearm_tx", "forearm_ty"),
                 physics_timestep: float = _PHYSICS_TIMESTEP, control_timestep: float = _CONTROL_TIMESTEP) -> None:
        del disable_colorization, change_color_on_activation, primitive_fingertip_collisions  # visual / meshes absent
        if not isinstance(hand_side, HandSide):
            raise ValueError("hand_side must be a HandSide")
        self._hand_side = hand_side
        self._side = "left" if hand_side == HandSide.LEFT else "right"
        self._reduced_action_space = bool(reduced_action_space)
        self._randomize_hand_positions = False
        # the reference's flag alone decides (:91): an unfingered MIDI gives every note finger 0
        self._disable_fingering_reward = bool(disable_fingering_reward)
        if n_seconds_lookahead is not None:
            n_steps_lookahead = int(np.ceil(n_seconds_lookahead / control_timestep))
        _PianoTaskBase.__init__(self, midi, n_steps_lookahead, trim_silence, physics_timestep, control_timestep,
                                initial_buffer_time=initial_buffer_time, augmentations=augmentations)
        self._disable_forearm_reward = True
        self._wrong_press_termination = wrong_press_termination
        self._disable_hand_collisions = False
        self._energy_penalty_coef = _ENERGY_PENALTY_COEF
        self.model = scene_compile.compile_scene(
            hands=True, gravity_compensation=gravity_compensation, physics_timestep=physics_timestep,
            reduced_action_space=self._reduced_action_space, attachment_yaw=float(attachment_yaw),
            forearm_dofs=tuple(forearm_dofs), absent_hand="right" if self._side == "left" else "left")
        names = ["key_press_reward", "sustain_reward", "energy_reward"]
        if not self._disable_fingering_reward:
            names.append("fingering_reward")
        self.reward_names = tuple(names)
        # `<hand>/position` = xpos of the hand's root body (hands/base.py:110-113): its compiled position plus the
        # two forearm slides along their world axes - affine in two columns of joints_pos
        a = self.model.a
        pre = "lh_" if self._side == "left" else "rh_"
        body = self.model.names["body"].index(f"{pre}shadow_hand/{pre}forearm")
        from ..scene.so3 import qrot
        quat = a["body_quat"][body]
        self._position_const = np.asarray(a["body_pos"][body], np.float64)
        self._position_axes = {}
        for name in ("forearm_tx", "forearm_ty"):
            d = self.model.names["dof"].index(f"{pre}shadow_hand/{name}")
            if d not in self.model.locked_dofs:
                self._position_axes[name] = qrot(quat, a["jnt_axis"][d])

    def _fingering_mode(self) -> int:
        return 2 if self._disable_fingering_reward else 1

    def _own(self, finger: int) -> bool:
        return finger >= 5 if self._side == "left" else finger < 5

    def _filter_tables(self, tables: SongTables) -> SongTables:
        """`_update_fingering_state` (:297-318): only this hand's (key, finger) pairs count; goal keeps every note."""
        pairs = np.full_like(tables.pairs, -1)
        for t in range(tables.n_frames):
            own = [(k, f) for k, f in tables.pairs[t] if k >= 0 and self._own(int(f))]
            for p, kf in enumerate(own):
                pairs[t, p] = kf
        return SongTables(goal=tables.goal, pairs=pairs, n_frames=tables.n_frames, pmax=tables.pmax)

    def _reset_trajectory(self) -> None:
        super()._reset_trajectory()
        self.song = self._filter_tables(self.song)

    def draw_song(self, random_state):
        midi, notes, sustains, tables = super().draw_song(random_state)
        return midi, notes, sustains, self._filter_tables(tables)

    def _extra_dropped_columns(self, full) -> set:
        if "fingering" not in full:
            return set()
        off, _ = full["fingering"]  # (2, 5) flags: right hand's row first
        return set(range(off + 5, off + 10)) if self._side == "right" else set(range(off, off + 5))

    def reduced_spec(self):
        act_keep, obs_keep, layout = super().reduced_spec()
        key = ("lh" if self._side == "left" else "rh") + "_shadow_hand/joints_pos"
        ordered = OrderedDict()
        for k, v in layout.items():  # composer order: entity observables (piano, hand), then goal, fingering
            ordered[k] = v
            if k == key:
                ordered[key.replace("joints_pos", "position")] = (len(obs_keep), 3)
        return act_keep, obs_keep, ordered

    def position_dofs(self):
        """Dof indices (engine state rows) of the forearm slides behind `position_map`'s columns, same order."""
        pre = "lh_" if self._side == "left" else "rh_"
        return [self.model.names["dof"].index(f"{pre}shadow_hand/{n}") for n in ("forearm_tx", "forearm_ty")
                if n in self._position_axes]

    def position_map(self):
        """(columns of the compacted observation row, W [k, 3], c [3]): position = c + row[columns] @ W."""
        _, _, layout = self.reduced_spec()
        off, size = layout[("lh" if self._side == "left" else "rh") + "_shadow_hand/joints_pos"]
        n_fore = len(self._position_axes)
        cols = [off + size - n_fore + i for i in range(n_fore)]  # hand.joints: the forearm dofs come last
        W = np.stack([self._position_axes[n] for n in ("forearm_tx", "forearm_ty") if n in self._position_axes]) \
            if n_fore else np.zeros((0, 3))
        return cols, W, self._position_const
