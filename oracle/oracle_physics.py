"""ctypes binding of the CPU fp64 oracle (oracle/piano_oracle.c). TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs - never by the product package. "parity unpinned" vs real mj_step
(see piano_oracle.h).
"""

from __future__ import annotations

import ctypes
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None

_PD = ctypes.POINTER(ctypes.c_double)


class _Contact(ctypes.Structure):
    _fields_ = [("geom1", ctypes.c_int), ("geom2", ctypes.c_int), ("dist", ctypes.c_double),
                ("pos", ctypes.c_double * 3), ("frame", ctypes.c_double * 9)]


class _Data(ctypes.Structure):
    _fields_ = (
        [(n, ctypes.c_int) for n in ("nv", "nbody", "ngeom", "nsite", "nu")]
        + [(n, _PD) for n in ("qpos", "qvel", "qacc_warmstart", "ctrl", "qfrc_applied", "xpos", "xquat",
                              "xmat", "xipos", "ximat", "xanchor", "xaxis", "geom_xpos", "geom_xmat",
                              "site_xpos", "qM", "qfrc_bias", "qfrc_passive", "actuator_force",
                              "actuator_velocity", "qfrc_actuator", "qfrc_smooth", "qacc_smooth")]
        + [("ncon", ctypes.c_int), ("nefc", ctypes.c_int), ("ncon_dropped", ctypes.c_int),
           ("contact", _Contact * 256), ("qacc", _PD), ("qfrc_constraint", _PD),
           ("solver_iter", ctypes.c_int), ("solver_cost", ctypes.c_double), ("ws", ctypes.c_void_p)]
    )


def build(force: bool = False) -> Path:
    so = _HERE / "libpiano_oracle.so"
    src = _HERE / "piano_oracle.c"
    if force or not so.exists() or so.stat().st_mtime < src.stat().st_mtime:
        subprocess.run(["make", "-C", str(_HERE)], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(str(build()))
        _LIB.po_data_create.restype = ctypes.POINTER(_Data)
        _LIB.po_data_create.argtypes = [ctypes.c_void_p]
        _LIB.po_data_free.argtypes = [ctypes.POINTER(_Data)]
        _LIB.po_reset.argtypes = [ctypes.c_void_p, ctypes.POINTER(_Data)]
        _LIB.po_step.argtypes = [ctypes.c_void_p, ctypes.POINTER(_Data)]
        _LIB.po_forward_position.argtypes = [ctypes.c_void_p, ctypes.POINTER(_Data), ctypes.c_int]
        _LIB.po_step_batch.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.POINTER(_Data)),
                                       ctypes.c_int, ctypes.c_int]
        _LIB.po_set_num_threads.argtypes = [ctypes.c_int]
        _LIB.po_step_sparse.argtypes = [ctypes.c_void_p, ctypes.POINTER(_Data)]
        _LIB.po_step_batch_sparse.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.POINTER(_Data)),
                                              ctypes.c_int, ctypes.c_int]
    return _LIB


class OraclePhysics:
    """One environment's physics state + the stage outputs of the last step."""

    _SHAPES = dict(qpos="nv", qvel="nv", qacc_warmstart="nv", ctrl="nu", qfrc_applied="nv",
                   xpos=("nbody", 3), xquat=("nbody", 4), xmat=("nbody", 9), xipos=("nbody", 3),
                   xanchor=("nv", 3), xaxis=("nv", 3), geom_xpos=("ngeom", 3), geom_xmat=("ngeom", 9),
                   site_xpos=("nsite", 3), qM=("nv", "nv"), qfrc_bias="nv", qfrc_passive="nv",
                   actuator_force="nu", actuator_velocity="nu", qfrc_actuator="nv", qfrc_smooth="nv",
                   qacc_smooth="nv", qacc="nv", qfrc_constraint="nv")

    def __init__(self, model):
        from diffusion_piano_b200.scene.compile import to_ctypes  # tables only, no compute

        self.model = model
        self._cm = to_ctypes(model)
        # private copy of body_pos: Entity.shift_pose (randomize_hand_positions) edits it per environment
        self._body_pos0 = np.array(model.a["body_pos"], np.float64, copy=True)
        self._body_pos = self._body_pos0.copy()
        self._cm.body_pos = self._body_pos.ctypes.data_as(type(self._cm.body_pos))
        self._mp = ctypes.cast(ctypes.pointer(self._cm), ctypes.c_void_p)
        self._d = lib().po_data_create(self._mp)
        d = self._d.contents
        dims = dict(nv=d.nv, nbody=d.nbody, ngeom=d.ngeom, nsite=d.nsite, nu=d.nu)
        for name, shape in self._SHAPES.items():
            shape = (shape,) if isinstance(shape, str) else shape
            shape = tuple(dims[s] if isinstance(s, str) else s for s in shape)
            n = int(np.prod(shape))
            if n == 0:
                arr = np.zeros(shape)
            else:
                arr = np.ctypeslib.as_array(getattr(d, name), shape=(n,)).reshape(shape)
            setattr(self, name, arr)

    def __del__(self):
        try:
            lib().po_data_free(self._d)
        except Exception:
            pass

    def reset(self):
        lib().po_reset(self._mp, self._d)

    def shift_hands(self, y_offset: float):
        """World-y shift of both hand root bodies relative to the compiled model (0 restores it)."""
        names = self.model.names["body"]
        for i, n in enumerate(names):
            if n.endswith("_forearm"):
                self._body_pos[i, 1] = self._body_pos0[i, 1] + float(y_offset)

    def step(self, nsub: int = 1):
        for _ in range(nsub):
            lib().po_step(self._mp, self._d)

    def step_sparse(self, nsub: int = 1):
        """po_step_sparse: the same step with the model's block structure exploited (the timed CPU baseline)."""
        for _ in range(nsub):
            lib().po_step_sparse(self._mp, self._d)

    def forward_position(self, with_collision: bool = True):
        lib().po_forward_position(self._mp, self._d, int(with_collision))

    @property
    def ncon(self):
        return self._d.contents.ncon

    @property
    def nefc(self):
        return self._d.contents.nefc

    @property
    def solver_iter(self):
        return self._d.contents.solver_iter

    def contacts(self):
        d = self._d.contents
        return [dict(geom1=c.geom1, geom2=c.geom2, dist=c.dist, pos=np.array(c.pos[:]),
                     frame=np.array(c.frame[:]).reshape(3, 3)) for c in d.contact[: d.ncon]]
