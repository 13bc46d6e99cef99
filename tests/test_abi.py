"""The C-ABI library loads and exports every symbol include/piano_step.h declares; without a GPU
pp_create must fail loudly (no CPU fallback)."""
import ctypes
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_all_declared_symbols(cuda_lib):
    header = (ROOT / "include" / "piano_step.h").read_text()
    declared = set(re.findall(r"\b(pp_[a-z_]+)\s*\(", header))
    from diffusion_piano_b200 import _lib
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(cuda_lib, name), name
    assert b"sm_100a" in cuda_lib.pp_version()


def test_model_struct_layout_matches_header():
    from diffusion_piano_b200.scene.compile import PPModelStruct
    header = (ROOT / "include" / "pp_model.h").read_text()
    start = header.index("typedef struct pp_model {") + len("typedef struct pp_model {")
    body = header[start:header.index("} pp_model;")]
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl or decl.startswith("typedef"):
            continue
        decl = re.sub(r"^(const\s+)?(int32_t|double)\s+", "", decl)
        for part in decl.split(","):
            names.append(re.sub(r"[\*\s]|\[.*", "", part))
    assert names == [n for n, _ in PPModelStruct._fields_]


def test_create_without_gpu_fails_loudly(cuda_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from diffusion_piano_b200 import _lib, music, suite
    task = suite.SelfActuatedPiano(music.load("CMajorScaleOneHand"))
    with pytest.raises(_lib.PianoStepError, match="no CPU fallback"):
        suite.BatchedEnvironment(task, 2)


def test_product_does_not_import_oracle():
    for path in (ROOT / "diffusion-piano_b200").rglob("*.py"):
        text = path.read_text()
        assert not re.search(r"^\s*(from|import)\s+\.*oracle|libpiano_oracle|oracle[/.]", text, re.M), path
    for path in (ROOT / "diffusion-piano_b200" / "csrc").iterdir():
        if path.suffix in (".cu", ".cuh"):
            assert "oracle/" not in path.read_text(), path


def test_ppo_library_exports_all_declared_symbols():
    """libppo_kernels.so (include/ppo_kernels.h): every declared entry point is exported and bound, and the
    control-block layout of the binding matches the library's."""
    header = (ROOT / "include" / "ppo_kernels.h").read_text()
    declared = set(re.findall(r"\b(ppo_[a-z_]+)\s*\(", header))
    from diffusion_piano_b200.ppo import _kernels
    assert declared == set(_kernels.EXPORTED_SYMBOLS)
    lib = _kernels.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.ppo_version()
    assert lib.ppo_ctrl_size() == ctypes.sizeof(_kernels.PPOCtrl)


def test_ppo_agent_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from diffusion_piano_b200.ppo import PPOAgent, _kernels
    with pytest.raises(_kernels.PPOKernelError, match="no CPU fallback"):
        PPOAgent(329, 45)
