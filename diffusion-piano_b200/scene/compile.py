"""Model compiler: scene description -> flat fp64 tables (`CompiledModel`).

This plays the role MuJoCo's XML compiler plays for the reference (`mjcf.Physics` behind
`composer_utils.Environment`, reference `parallelized_base_v2.py:39-43`): it resolves
defaults, lays bodies/DoFs/geoms/actuators out in MuJoCo's order, and precomputes the
compile-time constants the constraint solver needs (`dof_invweight0`, `body_invweight0`,
critical damping of the forearm slides, mixed contact parameters).

Layouts (two-hand scene):  bodies  0 world | 1..88 keys | 89..113 rh | 114..138 lh
                           dofs    0..87 keys | 88..113 rh | 114..139 lh
  per-hand dof order: forearm_tx, forearm_ty, WRJ2, WRJ1, FFJ4..1, MFJ4..1, RFJ4..1,
  LFJ5..1, THJ5..1 (MuJoCo qpos order: a body's own joints precede its children).
  `hand.joints` / `joints_pos` observable order is the MJCF document order in which
  dm_control finds the elements: the 24 menagerie joints first, forearm_tx/ty last
  (reference shadow_hand.py:159-184,272-311; shadow_hand_test.py:101-106).
"""

from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import numpy as np

from . import piano_model as pm
from . import shadow_hand_model as shm
from .so3 import (axis_angle, mirror_axis, mirror_quat, mirror_vec, qmat, qmul, qnorm, qrot)

GEOM_SPHERE, GEOM_CAPSULE, GEOM_BOX = 2, 3, 6
JNT_SLIDE, JNT_HINGE = 2, 3
CLS_KEY, CLS_BASE, CLS_RH_FINGER, CLS_LH_FINGER, CLS_RH_FOREARM, CLS_LH_FOREARM = range(6)
PAIRCLS_PIANO, PAIRCLS_HAND = 0, 1

DEFAULT_SOLREF = (0.02, 1.0)
DEFAULT_SOLIMP = (0.9, 0.95, 0.001, 0.5, 2.0)
DEFAULT_FRICTION = 1.0

# reference suite/tasks/base.py:34-37
RIGHT_HAND_POSITION = (0.4, 0.15, 0.13)
LEFT_HAND_POSITION = (0.4, -0.15, 0.13)
ABSENT_HAND_LIFT = 5.0  # metres: where the one-hand task parks the hand it does not have
HAND_QUATERNION = (-1.0, -1.0, 1.0, 1.0)


@dataclass
class CompiledModel:
    """Flat tables. All float arrays are float64, int arrays int32, C-contiguous."""

    opt: Dict[str, float]
    a: Dict[str, np.ndarray]  # the tables, by MuJoCo-style name
    names: Dict[str, List[str]] = field(default_factory=dict)
    has_hands: bool = True
    piano_actuators: bool = False

    @property
    def nv(self) -> int:
        return int(self.a["jnt_type"].shape[0])

    @property
    def nbody(self) -> int:
        return int(self.a["body_parent"].shape[0])

    @property
    def ngeom(self) -> int:
        return int(self.a["geom_type"].shape[0])

    @property
    def nu(self) -> int:
        return int(self.a["act_trn0"].shape[0])

    @property
    def nsite(self) -> int:
        return int(self.a["site_body"].shape[0])

    @property
    def npair(self) -> int:
        return int(self.a["pair_geom1"].shape[0])


class _Builder:
    def __init__(self):
        self.body = dict(parent=[0], pos=[(0, 0, 0)], quat=[(1, 0, 0, 0)], ipos=[(0, 0, 0)],
                         iquat=[(1, 0, 0, 0)], mass=[0.0], inertia=[(0, 0, 0)], gravcomp=[0.0],
                         dofadr=[-1], dofnum=[0], name=["world"])
        self.dof = dict(type=[], body=[], pos=[], axis=[], limited=[], range=[], stiffness=[],
                        springref=[], qpos0=[], damping=[], armature=[], frictionloss=[],
                        parent=[], name=[])
        self.geom = dict(type=[], body=[], pos=[], quat=[], size=[], cls=[], name=[])
        self.site = dict(body=[], pos=[], name=[])
        self.act = dict(trn0=[], trn1=[], kp=[], gain_only=[], ctrlrange=[], forcelimited=[],
                        forcerange=[], name=[])

    def add_body(self, name, parent, pos, quat, mass, ipos, iquat, inertia, gravcomp=0.0):
        b = self.body
        b["parent"].append(parent); b["pos"].append(tuple(pos)); b["quat"].append(tuple(qnorm(quat)))
        b["ipos"].append(tuple(ipos)); b["iquat"].append(tuple(qnorm(iquat)))
        b["mass"].append(mass); b["inertia"].append(tuple(inertia)); b["gravcomp"].append(gravcomp)
        b["dofadr"].append(-1); b["dofnum"].append(0); b["name"].append(name)
        return len(b["parent"]) - 1

    def add_dof(self, name, body, jtype, pos, axis, rng, stiffness=0.0, springref=0.0,
                damping=0.0, armature=0.0, frictionloss=0.0):
        d = self.dof
        idx = len(d["type"])
        b = self.body
        if b["dofnum"][body] == 0:
            b["dofadr"][body] = idx
            # parent dof = last dof of nearest ancestor that has dofs
            p = b["parent"][body]
            parent_dof = -1
            while p > 0:
                if b["dofnum"][p] > 0:
                    parent_dof = b["dofadr"][p] + b["dofnum"][p] - 1
                    break
                p = b["parent"][p]
        else:
            parent_dof = idx - 1
        b["dofnum"][body] += 1
        axis = np.asarray(axis, np.float64)
        axis = axis / np.linalg.norm(axis)
        d["type"].append(jtype); d["body"].append(body); d["pos"].append(tuple(pos))
        d["axis"].append(tuple(axis)); d["limited"].append(1 if rng is not None else 0)
        d["range"].append(tuple(rng) if rng is not None else (0.0, 0.0))
        d["stiffness"].append(stiffness); d["springref"].append(springref); d["qpos0"].append(0.0)
        d["damping"].append(damping); d["armature"].append(armature)
        d["frictionloss"].append(frictionloss); d["parent"].append(parent_dof); d["name"].append(name)
        return idx

    def add_geom(self, name, body, gtype, pos, quat, size, cls):
        g = self.geom
        size = tuple(size) + (0.0,) * (3 - len(size))
        g["type"].append(gtype); g["body"].append(body); g["pos"].append(tuple(pos))
        g["quat"].append(tuple(qnorm(quat))); g["size"].append(size); g["cls"].append(cls)
        g["name"].append(name)
        return len(g["type"]) - 1


def _add_piano(bld: _Builder, actuators: bool) -> None:
    for key in pm.build_keys():
        hx, hy, hz = key.half_size
        m = key.mass
        inertia = (m / 3 * (hy * hy + hz * hz), m / 3 * (hx * hx + hz * hz), m / 3 * (hx * hx + hy * hy))
        colour = "black" if key.black else "white"
        body = bld.add_body(f"piano/{colour}_key_{key.number}", 0, key.pos, (1, 0, 0, 0), m,
                            (0, 0, 0), (1, 0, 0, 0), inertia)
        dof = bld.add_dof(f"piano/{colour}_joint_{key.number}", body, JNT_HINGE, key.joint_pos,
                          (0, 1, 0), (0.0, key.max_angle), stiffness=pm.STIFFNESS,
                          springref=pm.SPRINGREF, damping=pm.DAMPING, armature=pm.ARMATURE)
        bld.add_geom(f"piano/{colour}_key_geom_{key.number}", body, GEOM_BOX, (0, 0, 0),
                     (1, 0, 0, 0), key.half_size, CLS_KEY)
        if actuators:  # piano_mjcf.py:56-62,80-81,99-100: general, gain 1, no bias
            a = bld.act
            a["trn0"].append(dof); a["trn1"].append(-1); a["kp"].append(1.0); a["gain_only"].append(1)
            a["ctrlrange"].append((0.0, key.max_angle)); a["forcelimited"].append(0)
            a["forcerange"].append((0.0, 0.0)); a["name"].append(f"piano/{colour}_actuator_{key.number}")
    bld.add_geom("piano/base", 0, GEOM_BOX, pm.BASE_POS, (1, 0, 0, 0), pm.BASE_SIZE, CLS_BASE)


def _add_hand(bld: _Builder, side: str, position, quaternion, gravcomp: bool, tx_range, remove_joints=()) -> dict:
    """Adds one hand; returns index bookkeeping for the task layer."""
    left = side == "left"
    prefix = "lh_" if left else "rh_"
    model = ("lh" if left else "rh") + "_shadow_hand/"
    mv = mirror_vec if left else (lambda v: np.asarray(v, np.float64))
    mq = mirror_quat if left else (lambda q: np.asarray(q, np.float64))
    ma = mirror_axis if left else (lambda a: np.asarray(a, np.float64))
    finger_cls = CLS_LH_FINGER if left else CLS_RH_FINGER
    forearm_cls = CLS_LH_FOREARM if left else CLS_RH_FOREARM
    info = dict(dof={}, body={}, geoms=[], dof_start=len(bld.dof["type"]), side=side)
    gc = 1.0 if gravcomp else 0.0

    def rec(node, parent):
        is_root = node["name"] == "forearm"
        pos = position if is_root else mv(node["pos"])
        quat = quaternion if is_root else mq(qnorm(node["quat"]))
        ine = node["inertial"]
        body = bld.add_body(model + prefix + node["name"], parent, pos, quat, ine["mass"],
                            mv(ine["pos"]), mq(qnorm(ine["quat"])), ine["diag"], gc)
        info["body"][node["name"]] = body
        if is_root:
            for dof_name, spec in shm.FOREARM_DOFS.items():
                rng = tx_range if dof_name == "forearm_tx" else spec["range"]
                # childclass defaults apply (armature, frictionloss); axis is NOT reflected
                # for tx/ty (shadow_hand.py:41-55 reflect=False). damping filled in later.
                info["dof"][dof_name] = bld.add_dof(
                    model + dof_name, body, JNT_SLIDE, (0, 0, 0), spec["axis"], rng,
                    damping=0.0, armature=shm.HAND_JOINT_DEFAULT["armature"],
                    frictionloss=shm.HAND_JOINT_DEFAULT["frictionloss"])
        elif node["joint"] is not None and node["joint"]["name"] not in remove_joints:
            j = dict(shm.HAND_JOINT_DEFAULT)
            j.update(shm.JOINT_CLASS[node["joint"]["cls"]])
            j.update({k: v for k, v in node["joint"].items() if k not in ("name", "cls")})  # per-joint overrides
            info["dof"][node["joint"]["name"]] = bld.add_dof(
                model + prefix + node["joint"]["name"], body, JNT_HINGE, (0, 0, 0), ma(j["axis"]),
                j["range"], damping=j["damping"], armature=j["armature"],
                frictionloss=j["frictionloss"])
        for gi, g in enumerate(node["geoms"]):
            gtype = GEOM_CAPSULE if g["type"] == "capsule" else GEOM_SPHERE
            gid = bld.add_geom(f"{model}{prefix}{node['name']}_c{gi}", body, gtype, mv(g["pos"]),
                               (1, 0, 0, 0), g["size"], forearm_cls if is_root else finger_cls)
            info["geoms"].append((node["name"], gid))
        if node["name"] in shm.FINGERTIP_BODIES:
            off = shm.THUMBTIP_OFFSET if node["name"] == "thdistal" else shm.FINGERTIP_OFFSET
            info.setdefault("site", {})[node["name"]] = (body, (0.0, 0.0, off))
        for child in node["children"]:
            rec(child, body)

    rec(shm.right_hand_tree(), 0)
    # fingertip sites in FINGERTIP_BODIES order (th, ff, mf, rf, lf)
    for name in shm.FINGERTIP_BODIES:
        body, pos = info["site"][name]
        bld.site["body"].append(body); bld.site["pos"].append(pos)
        bld.site["name"].append(f"{model}{name}_site")
    # actuators: 20 menagerie + forearm_tx + forearm_ty
    a = bld.act
    for name, kind, joints, cls in shm.ACTUATORS:
        if any(jn in remove_joints for jn in joints):
            continue  # the reference deletes the actuator with its joint (shadow_hand.py:164-171)
        c = shm.ACT_CLASS[cls]
        a["trn0"].append(info["dof"][joints[0]])
        a["trn1"].append(info["dof"][joints[1]] if kind == "tendon" else -1)
        a["kp"].append(c["kp"]); a["gain_only"].append(0); a["ctrlrange"].append(c["ctrlrange"])
        a["forcelimited"].append(1); a["forcerange"].append(c["forcerange"])
        a["name"].append(model + prefix + name)
    for dof_name, spec in shm.FOREARM_DOFS.items():
        rng = tx_range if dof_name == "forearm_tx" else spec["range"]
        a["trn0"].append(info["dof"][dof_name]); a["trn1"].append(-1); a["kp"].append(spec["stiffness"])
        a["gain_only"].append(0); a["ctrlrange"].append(tuple(rng)); a["forcelimited"].append(0)
        a["forcerange"].append((0.0, 0.0)); a["name"].append(model + dof_name)
    info["dof_end"] = len(bld.dof["type"])
    return info


# ------------------------------------------------------------------------------------
# compile-time dynamics at qpos0 (numpy, fp64): M0, dof_invweight0, body_invweight0
# ------------------------------------------------------------------------------------
def _kinematics0(bld: _Builder):
    """World poses at qpos0 (all joint displacements zero)."""
    nb = len(bld.body["parent"])
    xpos = np.zeros((nb, 3)); xquat = np.zeros((nb, 4)); xquat[0] = (1, 0, 0, 0)
    for i in range(1, nb):
        p = bld.body["parent"][i]
        xpos[i] = xpos[p] + qrot(xquat[p], bld.body["pos"][i])
        xquat[i] = qmul(xquat[p], np.asarray(bld.body["quat"][i]))
    return xpos, xquat


def _mass_matrix0(bld: _Builder):
    """M(qpos0) = sum_b Jp^T m Jp + Jr^T I Jr (+ armature), and per-body COM Jacobians."""
    nb = len(bld.body["parent"]); nv = len(bld.dof["type"])
    xpos, xquat = _kinematics0(bld)
    anchor = np.zeros((nv, 3)); axis = np.zeros((nv, 3))
    for d in range(nv):
        b = bld.dof["body"][d]
        anchor[d] = xpos[b] + qrot(xquat[b], bld.dof["pos"][d])
        axis[d] = qrot(xquat[b], bld.dof["axis"][d])
    M = np.zeros((nv, nv)); jacs = []
    for b in range(nb):
        jp = np.zeros((3, nv)); jr = np.zeros((3, nv))
        com = xpos[b] + qrot(xquat[b], bld.body["ipos"][b])
        # walk the dof chain from this body to the root
        bb = b
        while bb > 0:
            for k in range(bld.body["dofnum"][bb]):
                d = bld.body["dofadr"][bb] + k
                if bld.dof["type"][d] == JNT_HINGE:
                    jr[:, d] = axis[d]
                    jp[:, d] = np.cross(axis[d], com - anchor[d])
                else:
                    jp[:, d] = axis[d]
            bb = bld.body["parent"][bb]
        jacs.append((jp, jr))
        m = bld.body["mass"][b]
        if m > 0:
            R = qmat(qmul(xquat[b], np.asarray(bld.body["iquat"][b])))
            I = R @ np.diag(bld.body["inertia"][b]) @ R.T
            M += m * jp.T @ jp + jr.T @ I @ jr
    M += np.diag(bld.dof["armature"])
    return M, jacs


def _invweights(bld: _Builder, M, jacs):
    Minv = np.linalg.inv(M)
    nb = len(bld.body["parent"])
    dof_inv = np.diag(Minv).copy()
    body_inv = np.zeros((nb, 2))
    for b in range(1, nb):
        jp, jr = jacs[b]
        if not jp.any() and not jr.any():
            continue
        body_inv[b, 0] = np.trace(jp @ Minv @ jp.T) / 3
        body_inv[b, 1] = np.trace(jr @ Minv @ jr.T) / 3
    return dof_inv, body_inv


_FINGERS = ("ff", "mf", "rf", "lf")


def _hand_pairs(infos, disable_hand_collisions: bool):
    """Fixed candidate list for hand-hand colliders (see ASSUMPTIONS.md §collision)."""
    if disable_hand_collisions or not infos:
        return []
    pairs = []

    def of(info, pred):
        return [gid for name, gid in info["geoms"] if pred(name)]

    rh = next((i for i in infos if i["side"] == "right"), None)
    lh = next((i for i in infos if i["side"] == "left"), None)
    if rh and lh:
        pairs += [(a, b) for a in of(rh, lambda n: n == "forearm") for b in of(lh, lambda n: n == "forearm")]
    for info in infos:
        for f0, f1 in zip(_FINGERS[:-1], _FINGERS[1:]):
            seg0 = of(info, lambda n: n.startswith(f0) and n != "lfmetacarpal")
            seg1 = of(info, lambda n: n.startswith(f1) and n != "lfmetacarpal")
            pairs += [(a, b) for a in seg0 for b in seg1]
        thumb = of(info, lambda n: n in ("thmiddle", "thdistal"))
        for f in ("ff", "mf"):
            pairs += [(a, b) for a in thumb for b in of(info, lambda n: n.startswith(f))]
    tip = lambda n: n in ("thmiddle", "thdistal") or (n[2:] in ("middle", "distal") and n[:2] in _FINGERS)
    if rh and lh:
        pairs += [(a, b) for a in of(rh, tip) for b in of(lh, tip)]
    return pairs


def _mix(a, b):
    return tuple(0.5 * x + 0.5 * y for x, y in zip(a, b))


REDUCED_EXCLUDED_JOINTS = ("THJ5", "THJ1", "LFJ5")  # shadow_hand.py:73-77 (actuators A_<joint>)
REDUCED_THUMB_RANGE = (0.0, 0.698132)               # shadow_hand.py:79
LOCKED_ARMATURE = 1e8


def compile_scene(
    hands: bool = True,
    piano_actuators: bool = False,
    gravity_compensation: bool = False,
    disable_hand_collisions: bool = False,
    physics_timestep: float = 0.005,
    reduced_action_space: bool = False,
    reduced_by_removal: bool = False,
    attachment_yaw: float = 0.0,
    forearm_dofs=("forearm_tx", "forearm_ty"),
    absent_hand: Optional[str] = None,
    only_hand: Optional[str] = None,
) -> CompiledModel:
    """Builds the task scene the reference assembles in `suite/tasks/base.py:45-197`.

    reduced_action_space (`shadow_hand.py:73-79,160-181`): the reference deletes the joints THJ5, THJ1, LFJ5 of
    each hand together with their actuators (the bodies become welded at angle 0) and narrows THJ2 to
    (0, 0.698132). The step kernel has a fixed 26-dof tree per hand, so the joints are LOCKED instead: an armature
    of 1e8 kg m^2 (the dof's acceleration is below 1e-6 rad/s^2 for any force in the model and its coupling to the
    other dofs below float32 resolution), a range it never reaches, no frictionloss row, no damping, a dead actuator; the constants
    that MuJoCo derives from the mass matrix (meaninertia, invweight0) are computed on the matrix WITHOUT the
    locked rows, as the reference's compiled model would have them. `model.locked_dofs` / `model.dead_actuators`
    list them for the façade, which drops them from the action and observation specs.
    reduced_by_removal=True builds the reference's own form instead (joints and actuators deleted, nv = 134): the
    step kernel cannot run it, the CPU checker can: the tests compare the two forms.

    attachment_yaw [degrees] (`tasks/base.py:103,175-183`): the forearms are rotated about world z by +yaw (right
    hand) / -yaw (left hand) before the hand's own root quaternion.
    forearm_dofs (`tasks/base.py:104`, `shadow_hand.py:41-69,272-311`): which forearm joints exist. Subsets of the
    default pair are built by LOCKING the missing slide exactly as above (the kernel's tree keeps 26 dofs per hand);
    the reference's other four (forearm_tz / roll / pitch / yaw) would add dofs to that tree and are not built.

    absent_hand "left" / "right" (`piano_with_one_shadow_hand.py:96-102`: the other hand is detached from the
    scene): the step kernel always carries two 26-dof trees, so the absent hand stays in the tables but every one of
    its dofs is LOCKED as above, its actuators are dead and its root is parked ABSENT_HAND_LIFT metres above the
    keyboard where it touches nothing; the derived constants are those of the model without it.
    only_hand "left" / "right" builds the reference's own form instead (the other hand not in the model at all,
    nv = 114) for the CPU checker: the tests compare the two forms."""
    forearm_dofs = tuple(forearm_dofs)
    extra = [n for n in forearm_dofs if n not in shm.FOREARM_DOFS]
    if extra:
        raise NotImplementedError(f"forearm dofs {extra} are not built: the step kernel's hand tree has the two "
                                  "default slides only (forearm_tx, forearm_ty) - SURVEY 8f rank 4")
    bld = _Builder()
    _add_piano(bld, piano_actuators)
    infos = []
    if hands:
        half_w = pm.BASE_SIZE[1]  # piano.size[1], base.py:160
        for side, position in (("right", RIGHT_HAND_POSITION), ("left", LEFT_HAND_POSITION)):
            if only_hand is not None and side != only_hand:
                continue
            tx_range = (-half_w - position[1], half_w - position[1])  # from the hand's own position (base.py:160-164)
            if side == absent_hand:
                position = (position[0], position[1], position[2] + ABSENT_HAND_LIFT)
            sign = -1.0 if side == "left" else 1.0
            quat = qmul(axis_angle((0.0, 0.0, 1.0), np.radians(sign * attachment_yaw)), qnorm(HAND_QUATERNION))
            infos.append(_add_hand(bld, side, position, qnorm(quat), gravity_compensation, tx_range,
                                   REDUCED_EXCLUDED_JOINTS if reduced_action_space and reduced_by_removal else ()))
    locked, dead = [], []
    if hands and reduced_action_space and reduced_by_removal:  # the reference's own construction (for the CPU checker)
        for info in infos:
            pre = "lh_" if info["side"] == "left" else "rh_"
            d2 = next(i for i, n in enumerate(bld.dof["name"]) if n.endswith("/" + pre + "THJ2"))
            a2 = next(i for i, n in enumerate(bld.act["name"]) if n.endswith("/" + pre + "A_THJ2"))
            bld.dof["range"][d2] = REDUCED_THUMB_RANGE
            bld.act["ctrlrange"][a2] = REDUCED_THUMB_RANGE
    elif hands and reduced_action_space:
        for info in infos:
            pre = "lh_" if info["side"] == "left" else "rh_"
            for jn in REDUCED_EXCLUDED_JOINTS:
                locked.append(next(i for i, n in enumerate(bld.dof["name"]) if n.endswith("/" + pre + jn)))
                dead.append(next(i for i, n in enumerate(bld.act["name"]) if n.endswith("/" + pre + "A_" + jn)))
            d2 = next(i for i, n in enumerate(bld.dof["name"]) if n.endswith("/" + pre + "THJ2"))
            a2 = next(i for i, n in enumerate(bld.act["name"]) if n.endswith("/" + pre + "A_THJ2"))
            bld.dof["range"][d2] = REDUCED_THUMB_RANGE
            bld.act["ctrlrange"][a2] = REDUCED_THUMB_RANGE
    if hands and absent_hand is not None and only_hand is None:
        info = next(i for i in infos if i["side"] == absent_hand)
        model = ("lh" if absent_hand == "left" else "rh") + "_shadow_hand/"
        locked += list(range(info["dof_start"], info["dof_end"]))
        dead += [i for i, n in enumerate(bld.act["name"]) if n.startswith(model)]
    if hands and not reduced_by_removal and only_hand is None:
        for info in infos:  # forearm slides the task does not ask for: locked, their actuators dead
            model = "lh_shadow_hand/" if info["side"] == "left" else "rh_shadow_hand/"
            for dof_name in shm.FOREARM_DOFS:
                if dof_name not in forearm_dofs:
                    locked.append(info["dof"][dof_name])
                    dead.append(bld.act["name"].index(model + dof_name))
    M, jacs = _mass_matrix0(bld)
    if hands:
        # physics_utils.get_critical_damping_from_stiffness: 2*sqrt(dof_M0 * k)
        for info in infos:
            for dof_name, spec in shm.FOREARM_DOFS.items():
                d = info["dof"][dof_name]
                bld.dof["damping"][d] = 2.0 * np.sqrt(M[d, d] * spec["stiffness"])
    mean_inertia = float(np.mean(np.diag(M)))
    if locked:
        keep = np.setdiff1d(np.arange(M.shape[0]), locked)
        mean_inertia = float(np.mean(np.diag(M)[keep]))
        M_lock = M.copy()
        for d in locked:  # decouple and freeze: what deleting the joint does to every derived constant
            M_lock[d, :] = 0.0
            M_lock[:, d] = 0.0
            M_lock[d, d] = LOCKED_ARMATURE
            bld.dof["armature"][d] = LOCKED_ARMATURE
            bld.dof["damping"][d] = 0.0
            bld.dof["frictionloss"][d] = 0.0
            bld.dof["range"][d] = (-1.0, 1.0)  # the kernel wants every hand joint limited: a range it never reaches
        for a_ in dead:
            bld.act["kp"][a_] = 0.0
            bld.act["forcerange"][a_] = (0.0, 0.0)
        jacs_lock = []
        for jp, jr in jacs:
            jp, jr = jp.copy(), jr.copy()
            jp[:, locked] = 0.0
            jr[:, locked] = 0.0
            jacs_lock.append((jp, jr))
        dof_inv, body_inv = _invweights(bld, M_lock, jacs_lock)
    else:
        dof_inv, body_inv = _invweights(bld, M, jacs)
    pairs = _hand_pairs(infos, disable_hand_collisions)

    f64 = lambda x, shape=None: np.ascontiguousarray(np.asarray(x, np.float64).reshape(shape) if shape else np.asarray(x, np.float64))
    i32 = lambda x: np.ascontiguousarray(np.asarray(x, np.int32))
    nb = len(bld.body["parent"]); nv = len(bld.dof["type"]); ng = len(bld.geom["type"])
    a = dict(
        body_parent=i32(bld.body["parent"]), body_dofadr=i32(bld.body["dofadr"]),
        body_dofnum=i32(bld.body["dofnum"]), body_pos=f64(bld.body["pos"], (nb, 3)),
        body_quat=f64(bld.body["quat"], (nb, 4)), body_ipos=f64(bld.body["ipos"], (nb, 3)),
        body_iquat=f64(bld.body["iquat"], (nb, 4)), body_mass=f64(bld.body["mass"]),
        body_inertia=f64(bld.body["inertia"], (nb, 3)), body_gravcomp=f64(bld.body["gravcomp"]),
        body_invweight0=f64(body_inv, (nb, 2)),
        jnt_type=i32(bld.dof["type"]), jnt_body=i32(bld.dof["body"]), jnt_limited=i32(bld.dof["limited"]),
        dof_parent=i32(bld.dof["parent"]), jnt_pos=f64(bld.dof["pos"], (nv, 3)),
        jnt_axis=f64(bld.dof["axis"], (nv, 3)), jnt_range=f64(bld.dof["range"], (nv, 2)),
        jnt_stiffness=f64(bld.dof["stiffness"]), jnt_springref=f64(bld.dof["springref"]),
        qpos0=f64(bld.dof["qpos0"]), dof_damping=f64(bld.dof["damping"]),
        dof_armature=f64(bld.dof["armature"]), dof_frictionloss=f64(bld.dof["frictionloss"]),
        dof_invweight0=f64(dof_inv), dof_M0=f64(np.diag(M).copy()),
        geom_type=i32(bld.geom["type"]), geom_body=i32(bld.geom["body"]), geom_cls=i32(bld.geom["cls"]),
        geom_pos=f64(bld.geom["pos"], (ng, 3)), geom_quat=f64(bld.geom["quat"], (ng, 4)),
        geom_size=f64(bld.geom["size"], (ng, 3)),
        pair_geom1=i32([p[0] for p in pairs]), pair_geom2=i32([p[1] for p in pairs]),
        site_body=i32(bld.site["body"]), site_pos=f64(bld.site["pos"], (len(bld.site["body"]), 3)),
        act_trn0=i32(bld.act["trn0"]), act_trn1=i32(bld.act["trn1"]),
        act_gain_only=i32(bld.act["gain_only"]), act_forcelimited=i32(bld.act["forcelimited"]),
        act_kp=f64(bld.act["kp"]), act_ctrlrange=f64(bld.act["ctrlrange"], (len(bld.act["kp"]), 2)),
        act_forcerange=f64(bld.act["forcerange"], (len(bld.act["kp"]), 2)),
    )
    piano_solref = (2 * physics_timestep, 1.0)  # suite/tasks/base.py:66
    opt = dict(
        timestep=physics_timestep, gravity=(0.0, 0.0, -9.81),
        impratio=shm.OPTION_IMPRATIO if hands else 1.0, tolerance=1e-8, iterations=100,
        ls_iterations=50, meaninertia=mean_inertia,
        limit_solref=DEFAULT_SOLREF, limit_solimp=DEFAULT_SOLIMP,
        friction_solref=DEFAULT_SOLREF, friction_solimp=DEFAULT_SOLIMP,
        # equal priority, solmix 1/1 -> plain average; friction = max = 1; margin 0
        pair_solref=(_mix(shm.PLASTIC_SOLREF, piano_solref), shm.PLASTIC_SOLREF),
        pair_solimp=(_mix(shm.PLASTIC_SOLIMP, DEFAULT_SOLIMP), shm.PLASTIC_SOLIMP),
        pair_friction=(DEFAULT_FRICTION, DEFAULT_FRICTION), pair_margin=(0.0, 0.0),
    )
    names = dict(body=bld.body["name"], dof=bld.dof["name"], geom=bld.geom["name"],
                 site=bld.site["name"], actuator=bld.act["name"])
    model = CompiledModel(opt=opt, a=a, names=names, has_hands=hands, piano_actuators=piano_actuators)
    model.hand_info = infos
    model.locked_dofs, model.dead_actuators = sorted(locked), sorted(dead)
    return model


# ------------------------------------------------------------------------------------
# ctypes mirror of include/pp_model.h
# ------------------------------------------------------------------------------------
_PD = ctypes.POINTER(ctypes.c_double)
_PI = ctypes.POINTER(ctypes.c_int32)


class PPModelStruct(ctypes.Structure):
    _fields_ = (
        [(n, ctypes.c_int32) for n in ("nbody", "nv", "ngeom", "nsite", "nu", "npair", "nkey", "has_hands")]
        + [("timestep", ctypes.c_double), ("gravity", ctypes.c_double * 3),
           ("impratio", ctypes.c_double), ("tolerance", ctypes.c_double),
           ("iterations", ctypes.c_int32), ("ls_iterations", ctypes.c_int32),
           ("meaninertia", ctypes.c_double)]
        + [(n, _PI) for n in ("body_parent", "body_dofadr", "body_dofnum")]
        + [(n, _PD) for n in ("body_pos", "body_quat", "body_ipos", "body_iquat", "body_mass",
                              "body_inertia", "body_gravcomp", "body_invweight0")]
        + [(n, _PI) for n in ("jnt_type", "jnt_body", "jnt_limited", "dof_parent")]
        + [(n, _PD) for n in ("jnt_pos", "jnt_axis", "jnt_range", "jnt_stiffness", "jnt_springref",
                              "qpos0", "dof_damping", "dof_armature", "dof_frictionloss",
                              "dof_invweight0")]
        + [("limit_solref", ctypes.c_double * 2), ("limit_solimp", ctypes.c_double * 5),
           ("friction_solref", ctypes.c_double * 2), ("friction_solimp", ctypes.c_double * 5)]
        + [(n, _PI) for n in ("geom_type", "geom_body", "geom_cls")]
        + [(n, _PD) for n in ("geom_pos", "geom_quat", "geom_size")]
        + [("pair_solref", (ctypes.c_double * 2) * 2), ("pair_solimp", (ctypes.c_double * 5) * 2),
           ("pair_friction", ctypes.c_double * 2), ("pair_margin", ctypes.c_double * 2)]
        + [(n, _PI) for n in ("pair_geom1", "pair_geom2")]
        + [("site_body", _PI), ("site_pos", _PD)]
        + [(n, _PI) for n in ("act_trn0", "act_trn1", "act_gain_only", "act_forcelimited")]
        + [(n, _PD) for n in ("act_kp", "act_ctrlrange", "act_forcerange")]
    )


def to_ctypes(model: CompiledModel) -> PPModelStruct:
    """Returns a struct whose pointers alias model.a (keep `model` alive while it is used)."""
    s = PPModelStruct()
    s.nbody, s.nv, s.ngeom, s.nsite, s.nu, s.npair = (model.nbody, model.nv, model.ngeom,
                                                       model.nsite, model.nu, model.npair)
    s.nkey = pm.NUM_KEYS
    s.has_hands = int(model.has_hands)
    o = model.opt
    s.timestep = o["timestep"]; s.impratio = o["impratio"]; s.tolerance = o["tolerance"]
    s.iterations = int(o["iterations"]); s.ls_iterations = int(o["ls_iterations"])
    s.gravity[:] = o["gravity"]; s.meaninertia = o["meaninertia"]
    s.limit_solref[:] = o["limit_solref"]; s.limit_solimp[:] = o["limit_solimp"]
    s.friction_solref[:] = o["friction_solref"]; s.friction_solimp[:] = o["friction_solimp"]
    for c in range(2):
        s.pair_solref[c][:] = o["pair_solref"][c]; s.pair_solimp[c][:] = o["pair_solimp"][c]
    s.pair_friction[:] = o["pair_friction"]; s.pair_margin[:] = o["pair_margin"]
    for name, ctype in PPModelStruct._fields_:
        if ctype in (_PD, _PI):
            arr = model.a[name]
            setattr(s, name, arr.ctypes.data_as(ctype))
    s._keepalive = model
    return s
