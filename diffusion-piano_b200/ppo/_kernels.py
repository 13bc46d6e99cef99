"""ctypes binding of libppo_kernels.so (include/ppo_kernels.h). No fallback: a missing library raises."""

from __future__ import annotations

import ctypes
from pathlib import Path

import os

# PPO_KERNELS_LIB: an alternative build of the same library (A/B measurements of kernel variants)
LIB_PATH = Path(os.environ.get("PPO_KERNELS_LIB") or Path(__file__).resolve().parents[1] / "libppo_kernels.so")
_lib = None

P, I, F, U32, U64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint32, ctypes.c_uint64


class PPOCtrl(ctypes.Structure):
    """Host mirror of `ppo_ctrl` (read back for logging / checkpoints only)."""
    _fields_ = [("cursor", ctypes.c_int32), ("epoch", ctypes.c_int32), ("rng_counter", U64), ("seed", U64),
                ("lr", F * 2), ("adam_t", ctypes.c_int32 * 2), ("norm2", F * 2), ("best", F * 2),
                ("num_bad", ctypes.c_int32 * 2), ("actor_loss", F), ("entropy", F), ("mean_ratio", F),
                ("critic_loss", F), ("mean_value", F), ("mean_reward_n", F), ("pad_", ctypes.c_int32),
                ("rms_mean", ctypes.c_double), ("rms_var", ctypes.c_double), ("rms_count", ctypes.c_double)]


_SIGNATURES = {
    "ppo_last_error": (ctypes.c_char_p, []),
    "ppo_version": (ctypes.c_char_p, []),
    "ppo_ctrl_size": (I, []),
    "ppo_ctrl_init": (I, [P, U64, F, F, P]),
    "ppo_gemm_nt": (I, [P, P, P, I, I, I, P]),
    "ppo_gemm_nn": (I, [P, P, P, I, I, I, P]),
    "ppo_gemm_tn": (I, [P, P, P, I, I, I, P]),
    "ppo_relu_ln_fwd": (I, [P, P, P, P, P, P, P, I, I, F, P, P, P, U32, P]),
    "ppo_relu_ln_bwd_dy": (I, [P, P, P, P, P, F, P, I, I, P]),
    "ppo_ln_bwd_gemm_nn": (I, [P, P, P, P, P, F, P, P, P, I, I, I, P]),
    "ppo_relu_ln_colsums": (I, [P, P, P, P, F, P, P, P, P, I, I, P]),
    "ppo_relu_ln_bwd": (I, [P, P, P, P, P, F, P, P, P, P, I, I, P]),
    "ppo_policy_head": (I, [P, P, P, P, P, P, F, F, P, P, P, P, P, P, P, I, I, P]),
    "ppo_value_head_fused": (I, [P, P, P, P, P, P, P, P, I, I, P]),
    "ppo_value_head": (I, [P, P, P, P, P, P, I, P]),
    "ppo_add_bias": (I, [P, P, P, I, P]),
    "ppo_sample_actions": (I, [P, P, P, P, P, P, P, I, I, P]),
    "ppo_prepare_batch": (I, [P, P, P, P, F, F, P, P, P, P, P, I, P]),
    "ppo_gather_minibatch": (I, [P, P, P, P, P, P, I, P, U32, I, I, I, I, P, P, P, P, P, P]),
    "ppo_permutation": (I, [P, U32, I, P, P]),
    "ppo_grad_sumsq": (I, [P, I, P, P, P, I, P]),
    "ppo_grad_sumsq_static": (I, [P, I, P, I, P]),
    "ppo_clip_adam": (I, [P, P, P, P, I, P, I, F, F, F, F, F, P]),
    "ppo_advance": (I, [P, I, I, P]),
    "ppo_begin_update": (I, [P, P]),
    "ppo_plateau_step": (I, [P, F, I, P]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)


class PPOKernelError(RuntimeError):
    pass


def load():
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise PPOKernelError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                                 "g.build()'` (nvcc, sm_100a). There is no CPU fallback for the PPO update.")
        lib = ctypes.CDLL(str(LIB_PATH))
        for name, (restype, argtypes) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = restype, argtypes
        if lib.ppo_ctrl_size() != ctypes.sizeof(PPOCtrl):
            raise PPOKernelError(f"ppo_ctrl layout mismatch: library {lib.ppo_ctrl_size()} bytes, "
                                 f"binding {ctypes.sizeof(PPOCtrl)}")
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        raise PPOKernelError(f"ppo_kernels error {rc}: {load().ppo_last_error().decode()}")
