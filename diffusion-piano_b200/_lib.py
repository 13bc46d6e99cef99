"""ctypes binding of libpiano_step.so (include/piano_step.h).

The library is the product: if it is missing or cannot be loaded this module raises -
there is no CPU or PyTorch fallback for the step.
"""

from __future__ import annotations

import ctypes
from pathlib import Path

import os

_HERE = Path(__file__).resolve().parent
# PIANO_STEP_LIB: an alternative build of the same library (A/B measurements of kernel variants)
LIB_PATH = Path(os.environ.get("PIANO_STEP_LIB") or _HERE / "libpiano_step.so")
_lib = None

c_void_p, c_int, c_float_p = ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_float)


class PPConfig(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in (
        "num_envs", "device", "task", "n_substeps", "n_steps_lookahead", "n_frames", "pmax",
        "canonical_actions", "has_fingering", "disable_forearm_reward", "wrong_press_termination",
        "reward_type", "auto_reset", "max_contacts", "non_legacy_step")] + [
        ("energy_penalty_coef", ctypes.c_double), ("goal", c_void_p), ("pairs", c_void_p)]


TASK_SELF_ACTUATED_PIANO, TASK_PIANO_WITH_SHADOW_HANDS = 0, 1
REWARD_NEGATIVE_L2, REWARD_NEGATIVE_XENT = 0, 1
NUM_TERMS = 5
SCHED_AUTO, SCHED_LOCKSTEP, SCHED_DIVERGENT, SCHED_COMPACTED = 0, 1, 2, 3

_SIGNATURES = {
    "pp_create": (c_int, [c_void_p, ctypes.POINTER(PPConfig), ctypes.POINTER(c_void_p)]),
    "pp_destroy": (c_int, [c_void_p]),
    "pp_num_envs": (c_int, [c_void_p]),
    "pp_nv": (c_int, [c_void_p]),
    "pp_action_dim": (c_int, [c_void_p]),
    "pp_obs_dim": (c_int, [c_void_p]),
    "pp_action_bounds": (c_int, [c_void_p, c_void_p, c_void_p]),
    "pp_reset": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "pp_step": (c_int, [c_void_p] * 8),
    "pp_step_host": (c_int, [c_void_p] * 7),
    "pp_reset_host": (c_int, [c_void_p] * 3),
    "pp_host_timeline": (c_int, [c_void_p, c_int, c_void_p, c_void_p]),
    "pp_get_state": (c_int, [c_void_p] * 7),
    "pp_set_state": (c_int, [c_void_p] * 7),
    "pp_set_qfrc_applied": (c_int, [c_void_p] * 3),
    "pp_set_hand_offsets": (c_int, [c_void_p] * 3),
    "pp_set_song_bank": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int]),
    "pp_set_song_ids": (c_int, [c_void_p] * 3),
    "pp_set_substep_activations": (c_int, [c_void_p] * 2),
    "pp_get_metrics": (c_int, [c_void_p] * 4),
    "pp_get_diag": (c_int, [c_void_p] * 8),
    "pp_launch_count": (ctypes.c_longlong, [c_void_p]),
    "pp_set_schedule": (c_int, [c_void_p, c_int]),
    "pp_get_schedule": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "pp_schedule_in_force": (c_int, [c_void_p]),
    "pp_last_error": (ctypes.c_char_p, []),
    "pp_version": (ctypes.c_char_p, []),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class PianoStepError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise PianoStepError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). There is no CPU fallback for the step.")
        lib = ctypes.CDLL(str(LIB_PATH))
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = restype, argtypes
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise PianoStepError(f"piano_step error {rc}: {load().pp_last_error().decode()}")
