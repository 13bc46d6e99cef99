"""BatchedPianoEngine: thin Python owner of one pp_env handle (one per GPU).

torch is used only for device buffers and streams; every step goes through the C-ABI
(`pp_step` with device pointers, `pp_step_host` with numpy buffers).
"""

from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np

from . import _lib
from .music.midi_file import SongTables
from .scene.compile import CompiledModel, to_ctypes


@dataclass
class EngineConfig:
    num_envs: int
    task: int
    n_substeps: int
    n_steps_lookahead: int = 0
    canonical_actions: bool = False
    has_fingering: int = 0  # 1: fingering reward + observation, 0: optimal-transport fingering reward, 2: neither
    disable_forearm_reward: bool = False
    wrong_press_termination: bool = False
    reward_type: int = _lib.REWARD_NEGATIVE_L2
    auto_reset: bool = True
    max_contacts: int = 0
    legacy_step: bool = True  # dm_control Physics.legacy_step (False: derived quantities are one substep old)
    energy_penalty_coef: float = 5e-3
    device: int = 0


def _ptr(t) -> Optional[int]:
    """Device/host pointer of a torch tensor or numpy array (None passes NULL)."""
    if t is None:
        return None
    if isinstance(t, np.ndarray):
        assert t.flags["C_CONTIGUOUS"]
        return t.ctypes.data
    assert t.is_contiguous()
    return t.data_ptr()


class BatchedPianoEngine:
    def __init__(self, model: CompiledModel, song: SongTables, cfg: EngineConfig):
        self.lib = _lib.load()
        self.model, self.song, self.cfg = model, song, cfg
        self._cm = to_ctypes(model)
        self._goal = np.ascontiguousarray(song.goal, np.uint8)
        self._pairs = np.ascontiguousarray(song.pairs, np.int16)
        c = _lib.PPConfig()
        c.num_envs, c.device, c.task = cfg.num_envs, cfg.device, cfg.task
        c.n_substeps, c.n_steps_lookahead = cfg.n_substeps, cfg.n_steps_lookahead
        c.n_frames, c.pmax = song.n_frames, song.pmax
        c.canonical_actions, c.has_fingering = int(cfg.canonical_actions), int(cfg.has_fingering)
        c.disable_forearm_reward = int(cfg.disable_forearm_reward)
        c.wrong_press_termination = int(cfg.wrong_press_termination)
        c.reward_type, c.auto_reset, c.max_contacts = cfg.reward_type, int(cfg.auto_reset), cfg.max_contacts
        c.non_legacy_step = int(not cfg.legacy_step)
        c.energy_penalty_coef = cfg.energy_penalty_coef
        c.goal, c.pairs = self._goal.ctypes.data, self._pairs.ctypes.data
        h = ctypes.c_void_p()
        _lib.check(self.lib.pp_create(ctypes.cast(ctypes.pointer(self._cm), ctypes.c_void_p),
                                      ctypes.byref(c), ctypes.byref(h)))
        self._h = h
        self.num_envs = cfg.num_envs
        self.nv = self.lib.pp_nv(h)
        self.action_dim = self.lib.pp_action_dim(h)
        self.obs_dim = self.lib.pp_obs_dim(h)
        lo = np.zeros(self.action_dim, np.float32)
        hi = np.zeros(self.action_dim, np.float32)
        _lib.check(self.lib.pp_action_bounds(h, lo.ctypes.data, hi.ctypes.data))
        self.action_minimum, self.action_maximum = lo, hi

    def close(self):
        if getattr(self, "_h", None):
            self.lib.pp_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- device-pointer path (torch CUDA tensors)
    def reset(self, obs=None, mask=None, stream: int = 0):
        _lib.check(self.lib.pp_reset(self._h, _ptr(mask), _ptr(obs), stream))

    def step(self, actions, obs, reward, done, reward_terms=None, discount=None, stream: int = 0):
        _lib.check(self.lib.pp_step(self._h, _ptr(actions), _ptr(obs), _ptr(reward), _ptr(reward_terms),
                                    _ptr(done), _ptr(discount), stream))

    # ---- host-buffer path (numpy), copies inside the call
    def reset_host(self, obs: Optional[np.ndarray] = None, mask: Optional[np.ndarray] = None):
        _lib.check(self.lib.pp_reset_host(self._h, _ptr(mask), _ptr(obs)))

    def step_host(self, actions: np.ndarray, obs, reward, done, reward_terms=None, discount=None):
        _lib.check(self.lib.pp_step_host(self._h, _ptr(actions), _ptr(obs), _ptr(reward),
                                         _ptr(reward_terms), _ptr(done), _ptr(discount)))

    def host_timeline(self, enable: bool = True):
        """CUDA-event times [ms from the call's start] of the last step_host: (actions on the device,
        [(step done, results on the host) per piece]); arms / disarms the recording for later calls."""
        ms = (ctypes.c_float * 17)()
        n = ctypes.c_int(0)
        _lib.check(self.lib.pp_host_timeline(self._h, int(enable), ctypes.addressof(ms), ctypes.addressof(n)))
        return float(ms[0]), [(float(ms[1 + 2 * k]), float(ms[2 + 2 * k])) for k in range(n.value)]

    # ---- state / diagnostics (torch CUDA tensors)
    def get_state(self, qpos=None, qvel=None, warm=None, t_idx=None, sustain=None, stream: int = 0):
        _lib.check(self.lib.pp_get_state(self._h, _ptr(qpos), _ptr(qvel), _ptr(warm), _ptr(t_idx),
                                         _ptr(sustain), stream))

    def set_state(self, qpos=None, qvel=None, warm=None, t_idx=None, sustain=None, stream: int = 0):
        _lib.check(self.lib.pp_set_state(self._h, _ptr(qpos), _ptr(qvel), _ptr(warm), _ptr(t_idx),
                                         _ptr(sustain), stream))

    def set_qfrc_applied(self, qfrc, stream: int = 0):
        _lib.check(self.lib.pp_set_qfrc_applied(self._h, _ptr(qfrc), stream))

    def set_hand_offsets(self, y_off, stream: int = 0):
        """y shift [N] of both hands that each env takes at its next reset (None: zero again)."""
        _lib.check(self.lib.pp_set_hand_offsets(self._h, _ptr(y_off), stream))

    def set_song_bank(self, songs) -> None:
        """songs: sequence of `SongTables` (MIDI augmentations): uploaded back to back; every env plays the one
        whose id `set_song_ids` hands it before its reset. Synchronises the device; reset afterwards."""
        pmax = max(t.pmax for t in songs)
        goal = np.ascontiguousarray(np.concatenate([t.goal for t in songs], 0), np.uint8)
        pairs = np.full((goal.shape[0], pmax, 2), -1, np.int16)
        off = 0
        for t in songs:
            pairs[off:off + t.n_frames, :t.pmax] = t.pairs
            off += t.n_frames
        frames = np.ascontiguousarray([t.n_frames for t in songs], np.int32)
        self._bank = (goal, pairs, frames)
        _lib.check(self.lib.pp_set_song_bank(self._h, goal.ctypes.data, pairs.ctypes.data, frames.ctypes.data,
                                             len(songs), pmax))

    def set_song_ids(self, ids, stream: int = 0) -> None:
        """ids: int32 CUDA tensor [N] - the bank entry each env takes at its next reset."""
        _lib.check(self.lib.pp_set_song_ids(self._h, _ptr(ids), stream))

    def set_substep_activations(self, buf):
        """uint32 CUDA tensor [N, n_substeps, 4] that every later step fills with the key / pedal activation bits
        after each substep (None: stop). The caller keeps the tensor alive."""
        _lib.check(self.lib.pp_set_substep_activations(self._h, _ptr(buf)))

    def get_metrics(self, episode_metrics=None, episode_count=None, stream: int = 0):
        """Musical metrics of every env's last finished episode: [N,6] f32, and [N] i32 episode counts."""
        _lib.check(self.lib.pp_get_metrics(self._h, _ptr(episode_metrics), _ptr(episode_count), stream))

    def get_diag(self, qacc=None, actuator_force=None, actuator_velocity=None, site_xpos=None,
                 activation=None, counters=None, stream: int = 0):
        _lib.check(self.lib.pp_get_diag(self._h, _ptr(qacc), _ptr(actuator_force), _ptr(actuator_velocity),
                                        _ptr(site_xpos), _ptr(activation), _ptr(counters), stream))

    def set_schedule(self, mode) -> None:
        """Launch shape of the two-hands step: "auto" (default, timed on steps 4..15), "lockstep" (2 envs per
        block: every env replays the same actions), "divergent" (10 envs per block, ordered by expected work:
        actions differ per env) or "compacted" (prologue / Newton / finish launches per substep, Newton warps pulling
        envs from cost-class lists: large batches with per-env actions). Never changes results (include/piano_step.h)."""
        modes = {"auto": _lib.SCHED_AUTO, "lockstep": _lib.SCHED_LOCKSTEP, "divergent": _lib.SCHED_DIVERGENT,
                 "compacted": _lib.SCHED_COMPACTED}
        _lib.check(self.lib.pp_set_schedule(self._h, modes.get(mode, mode)))

    def get_schedule(self) -> Dict[str, int]:
        m, b, s = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        _lib.check(self.lib.pp_get_schedule(self._h, ctypes.addressof(m), ctypes.addressof(b), ctypes.addressof(s)))
        names = ("auto", "lockstep", "divergent", "compacted")
        return {"mode": names[m.value], "block_envs": b.value, "sorted_by_work": s.value,
                "in_force": names[self.lib.pp_schedule_in_force(self._h)]}

    @property
    def launch_count(self) -> int:
        return int(self.lib.pp_launch_count(self._h))
