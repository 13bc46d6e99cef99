"""ctypes binding of tools/experiments/libppo_experiments.so (product PPO kernels + the measured-slower experiments)."""
import ctypes
import subprocess
from pathlib import Path

HERE = Path(__file__).resolve().parent
P, I, F, U32 = ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_uint32
EXPERIMENTS = {
    "ppo_linear_relu_ln_fwd": (I, [P, P, P, P, P, P, P, P, I, I, I, F, P, P, P, U32, P]),
    "ppo_ln_gemm_nt": (I, [P, P, P, P, P, F, P, P, P, P, P, I, I, I, P]),
    "ppo_dropout_mask": (I, [P, I, I, F, P, U32, P]),
    "ppo_policy_head_fused": (I, [P, P, P, P, P, P, P, F, F, P, P, P, P, P, P, P, P, I, I, I, P]),
    "ppo_pdl_probe": (I, [P, I, I, P]),
    "ppo_allreduce_sumsq": (I, [P, P, P, I, I, I, P, P, P, P, P, I, P]),
    "ppo_allreduce_flag_words": (I, []),
}


def load_experiments():
    so = HERE / "libppo_experiments.so"
    if not so.exists():
        subprocess.run(["bash", str(HERE / "build.sh")], check=True)
    import sys
    sys.path.insert(0, str(HERE.parents[1]))
    from diffusion_piano_b200.ppo import _kernels
    lib = ctypes.CDLL(str(so))
    for name, (restype, argtypes) in list(_kernels._SIGNATURES.items()) + list(EXPERIMENTS.items()):
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = restype, argtypes
    return lib
