"""Shadow Hand E3M5 description used by the batched step.

PROVENANCE - READ THIS. The reference loads `right_hand.xml` / `left_hand.xml` from
mujoco_menagerie@1afc8be (reference `shadow_hand_constants.py:52-53`,
`scripts/install_deps.sh:78-88`). That submodule is EMPTY in the reference checkout and
MuJoCo is not installable here, so the numbers below are a RE-SPECIFICATION of the public
E3M5 model written from memory of the menagerie file. Structure (body tree, joint names and
order, actuator names and order, tendon coupling of the J0 actuators, default classes) is
what the reference code relies on (`shadow_hand.py:41-85,128-311`, `shadow_hand_test.py:81-124`);
numeric values (link offsets, inertias, gains, ranges, collision primitive sizes) are
assumptions and are listed in `scene/ASSUMPTIONS.md`. Hand-side parity versus real
`mj_step` is therefore UNPINNED; parity of the CUDA step is against the fp64 oracle that
consumes exactly these tables.

What the reference adds on top of the menagerie file and is reproduced here exactly:
* forearm DoFs `forearm_tx` (slide, axis (-1,0,0), kp 300) and `forearm_ty` (slide, axis
  (0,0,1), kp 300, range (0,0.06)), critically damped, with position actuators appended
  after the 20 menagerie actuators (`shadow_hand.py:41-55,272-311`);
* fingertip sites at z=0.026 (0.0275 thumb) on th/ff/mf/rf/lf distal (`:81-82,190-207`);
* root pose and forearm_tx range (`suite/tasks/base.py:34-37,160-194`).

The left hand is the mirror image of the right through the forearm-local x=0 plane:
positions x -> -x, quaternions (w,x,y,z) -> (w,x,-y,-z), joint axes (ax,ay,az) ->
(ax,-ay,-az) (so that equal joint angles give mirrored poses).

Collision geometry: the distal phalanx colliders are convex meshes in menagerie
(`*_distal_pst.obj`, not available) - replaced by fitted capsules, i.e. the reference's own
`primitive_fingertip_collisions=True` variant (`shadow_hand.py:144-152`). Only capsule and
sphere colliders of the fingers plus one forearm capsule are kept (palm/wrist boxes and
knuckle cylinders cannot reach the keys at the task's hand height; see ASSUMPTIONS.md).
"""

from __future__ import annotations

from typing import Dict, List

# ---- joint default classes (menagerie <default> tree, recollected) -----------------
HAND_JOINT_DEFAULT = dict(axis=(1.0, 0.0, 0.0), damping=0.05, armature=0.0002, frictionloss=0.01)
JOINT_CLASS: Dict[str, dict] = {
    "wrist_y": dict(axis=(0.0, 1.0, 0.0), range=(-0.523599, 0.174533), damping=0.5),
    "wrist_x": dict(range=(-0.698132, 0.488692), damping=0.5),
    "thbase": dict(axis=(0.0, 0.0, -1.0), range=(-1.0472, 1.0472)),
    "thproximal": dict(range=(0.0, 1.22173)),
    "thhub": dict(range=(-0.20944, 0.20944)),
    "thmiddle": dict(axis=(0.0, -1.0, 0.0), range=(-0.698132, 0.698132)),
    "thdistal": dict(range=(-0.261799, 1.5708)),
    "metacarpal": dict(axis=(0.573576, 0.0, 0.819152), range=(0.0, 0.785398)),
    "knuckle": dict(axis=(0.0, -1.0, 0.0), range=(-0.349066, 0.349066)),
    "proximal": dict(range=(-0.261799, 1.5708)),
    "middle_distal": dict(range=(0.0, 1.5708)),
}
# position actuator classes: kp, ctrlrange, forcerange
ACT_CLASS: Dict[str, dict] = {
    "wrist_y": dict(kp=10.0, ctrlrange=(-0.523599, 0.174533), forcerange=(-10.0, 10.0)),
    "wrist_x": dict(kp=8.0, ctrlrange=(-0.698132, 0.488692), forcerange=(-5.0, 5.0)),
    "thbase": dict(kp=0.4, ctrlrange=(-1.0472, 1.0472), forcerange=(-3.0, 3.0)),
    "thproximal": dict(kp=1.0, ctrlrange=(0.0, 1.22173), forcerange=(-2.0, 2.0)),
    "thhub": dict(kp=0.5, ctrlrange=(-0.20944, 0.20944), forcerange=(-1.0, 1.0)),
    "thmiddle": dict(kp=1.5, ctrlrange=(-0.698132, 0.698132), forcerange=(-1.0, 1.0)),
    "thdistal": dict(kp=1.0, ctrlrange=(-0.261799, 1.5708), forcerange=(-1.0, 1.0)),
    "metacarpal": dict(kp=1.0, ctrlrange=(0.0, 0.785398), forcerange=(-1.0, 1.0)),
    "knuckle": dict(kp=1.0, ctrlrange=(-0.349066, 0.349066), forcerange=(-1.0, 1.0)),
    "proximal": dict(kp=1.0, ctrlrange=(-0.261799, 1.5708), forcerange=(-1.0, 1.0)),
    "middle_distal": dict(kp=0.5, ctrlrange=(0.0, 3.1415), forcerange=(-1.0, 1.0)),
}
# plastic default class: contact softness of every hand collider
PLASTIC_SOLREF = (0.005, 1.0)
PLASTIC_SOLIMP = (0.5, 0.99, 0.0001, 0.5, 2.0)
OPTION_IMPRATIO = 10.0  # <option impratio="10"/> merges into the scene on attach

FINGERTIP_OFFSET = 0.026
THUMBTIP_OFFSET = 0.0275
FINGERTIP_BODIES = ("thdistal", "ffdistal", "mfdistal", "rfdistal", "lfdistal")

_I = (1.0, 0.0, 0.0, 0.0)
_Z90 = (1.0, 0.0, 0.0, 1.0)  # unnormalised, MuJoCo normalises


# Fingertip colliders. The menagerie distal colliders are convex meshes (f_distal_pst / th_distal_pst, absent);
# a capsule stands in for each. Its size is CALIBRATED, not recollected: of the sizes tried, these make the
# reference's recorded policy (examples/twinkle_twinkle_actions.npy, trained on the real meshes) press the most
# goal keys in the replay (tools/calibrate_search.py, scene/ASSUMPTIONS.md section 3): key-press F1 0.365 with the
# first guess (r 7.5 mm, tip at 29 mm / thumb r 9 mm, tip at 34 mm) -> 0.60 with these.
FINGER_DISTAL_CAPSULE = dict(type="capsule", size=(0.0095, 0.010), pos=(0, 0, 0.015))   # tip at z = 34.5 mm
THUMB_DISTAL_CAPSULE = dict(type="capsule", size=(0.013, 0.009), pos=(0, 0, 0.016))     # tip at z = 38 mm


def _finger(prefix: str, knuckle_pos, j4: str = "J4", j4_axis=None) -> dict:
    """knuckle -> proximal -> middle -> distal chain shared by ff/mf/rf/lf.
    j4_axis: the ring and little finger knuckles override the class axis (0,-1,0) with (0,1,0), as the Shadow
    Hand URDF does (positive J4 spreads the fingers away from the middle finger on both sides of it); the
    replay probe agrees (left ring finger 24 -> 34 of 43 presses, left little finger 2 -> 15 of 20)."""
    p = prefix
    j4_joint = dict(name=p.upper() + j4, cls="knuckle")
    if j4_axis is not None:
        j4_joint["axis"] = j4_axis
    return dict(
        name=p + "knuckle", pos=knuckle_pos, quat=_I,
        inertial=dict(mass=0.008, pos=(0, 0, 0), quat=(0.5, 0.5, -0.5, 0.5), diag=(3.2e-07, 2.6e-07, 2.6e-07)),
        joint=j4_joint, geoms=[],
        children=[dict(
            name=p + "proximal", pos=(0, 0, 0), quat=_I,
            inertial=dict(mass=0.03, pos=(0, 0, 0.0225), quat=_Z90, diag=(1e-05, 9.8e-06, 1.8e-06)),
            joint=dict(name=p.upper() + "J3", cls="proximal"),
            geoms=[dict(type="capsule", size=(0.009, 0.02), pos=(0, 0, 0.025))],
            children=[dict(
                name=p + "middle", pos=(0, 0, 0.045), quat=_I,
                inertial=dict(mass=0.017, pos=(0, 0, 0.0125), quat=_Z90, diag=(2.7e-06, 2.6e-06, 8.7e-07)),
                joint=dict(name=p.upper() + "J2", cls="middle_distal"),
                geoms=[dict(type="capsule", size=(0.009, 0.0125), pos=(0, 0, 0.0125))],
                children=[dict(
                    name=p + "distal", pos=(0, 0, 0.025), quat=_I,
                    inertial=dict(mass=0.013, pos=(0, 0, 0.0130769), quat=_Z90,
                                  diag=(1.28092e-06, 1.12092e-06, 5.3e-07)),
                    joint=dict(name=p.upper() + "J1", cls="middle_distal"),
                    geoms=[dict(FINGER_DISTAL_CAPSULE)],  # stands in for mesh f_distal_pst
                    children=[],
                )],
            )],
        )],
    )


def right_hand_tree() -> dict:
    """Body tree of the right hand in the forearm body's frame (names without rh_ prefix)."""
    lf = dict(
        name="lfmetacarpal", pos=(-0.033, 0, 0.02071), quat=_I,
        inertial=dict(mass=0.03, pos=(0, 0, 0.04), quat=_Z90, diag=(1.638e-05, 1.45e-05, 4.272e-06)),
        joint=dict(name="LFJ5", cls="metacarpal"), geoms=[],
        children=[_finger("lf", (0, 0, 0.06579), j4_axis=(0.0, 1.0, 0.0))],
    )
    thumb = dict(
        name="thbase", pos=(0.034, -0.00858, 0.029), quat=(0.92388, 0, 0.382683, 0),
        inertial=dict(mass=0.01, pos=(0, 0, 0), quat=_I, diag=(1.6e-07, 1.6e-07, 1.6e-07)),
        joint=dict(name="THJ5", cls="thbase"),
        geoms=[dict(type="sphere", size=(0.013,), pos=(0, 0, 0))],
        children=[dict(
            name="thproximal", pos=(0, 0, 0), quat=_I,
            inertial=dict(mass=0.04, pos=(0, 0, 0.019), quat=_I, diag=(1.36e-05, 1.36e-05, 3.13e-06)),
            joint=dict(name="THJ4", cls="thproximal"),
            geoms=[dict(type="capsule", size=(0.0105, 0.009), pos=(0, 0, 0.02))],
            children=[dict(
                name="thhub", pos=(0, 0, 0.038), quat=_I,
                inertial=dict(mass=0.005, pos=(0, 0, 0), quat=_I, diag=(1e-06, 1e-06, 3e-07)),
                joint=dict(name="THJ3", cls="thhub"),
                geoms=[dict(type="sphere", size=(0.011,), pos=(0, 0, 0))],
                children=[dict(
                    name="thmiddle", pos=(0, 0, 0), quat=_I,
                    inertial=dict(mass=0.02, pos=(0, 0, 0.016), quat=_I, diag=(5.1e-06, 5.1e-06, 1.21e-06)),
                    joint=dict(name="THJ2", cls="thmiddle"),
                    geoms=[dict(type="capsule", size=(0.009, 0.009), pos=(0, 0, 0.012)),
                           dict(type="sphere", size=(0.01,), pos=(0, 0, 0.03))],
                    children=[dict(
                        name="thdistal", pos=(0, 0, 0.032), quat=(1.0, 0.0, 0.0, -1.0),
                        inertial=dict(mass=0.017, pos=(0, 0, 0.0145588), quat=_Z90,
                                      diag=(2.37794e-06, 2.27794e-06, 1e-06)),
                        joint=dict(name="THJ1", cls="thdistal"),
                        geoms=[dict(THUMB_DISTAL_CAPSULE)],  # stands in for mesh th_distal_pst
                        children=[],
                    )],
                )],
            )],
        )],
    )
    palm = dict(
        name="palm", pos=(0, 0, 0.034), quat=_I,
        inertial=dict(mass=0.3, pos=(0, 0, 0.035), quat=_Z90, diag=(0.0005287, 0.0003581, 0.000191)),
        joint=dict(name="WRJ1", cls="wrist_x"), geoms=[],
        children=[_finger("ff", (0.033, 0, 0.095)), _finger("mf", (0.011, 0, 0.099)),
                  _finger("rf", (-0.011, 0, 0.095), j4_axis=(0.0, 1.0, 0.0)), lf, thumb],
    )
    wrist = dict(
        name="wrist", pos=(0.01, 0, 0.21301), quat=_I,
        inertial=dict(mass=0.1, pos=(0, 0, 0.029), quat=(0.5, 0.5, 0.5, 0.5), diag=(6.4e-05, 4.38e-05, 3.5e-05)),
        joint=dict(name="WRJ2", cls="wrist_y"), geoms=[], children=[palm],
    )
    forearm = dict(
        name="forearm", pos=(0, 0, 0), quat=_I,  # pose is overwritten by the task
        inertial=dict(mass=3.0, pos=(0, 0, 0.09), quat=_I, diag=(0.0138, 0.0138, 0.00744)),
        joint=None,  # forearm_tx / forearm_ty are added by the compiler
        # stands in for mesh forearm_collision (+ the wrist-end box): one capsule
        geoms=[dict(type="capsule", size=(0.045, 0.085), pos=(0, 0, 0.085))],
        children=[wrist],
    )
    return forearm


# Menagerie actuator order (then the reference appends forearm_tx, forearm_ty).
# (name, transmission kind, target joints, class)
ACTUATORS: List[tuple] = [
    ("A_WRJ2", "joint", ("WRJ2",), "wrist_y"),
    ("A_WRJ1", "joint", ("WRJ1",), "wrist_x"),
    ("A_THJ5", "joint", ("THJ5",), "thbase"),
    ("A_THJ4", "joint", ("THJ4",), "thproximal"),
    ("A_THJ3", "joint", ("THJ3",), "thhub"),
    ("A_THJ2", "joint", ("THJ2",), "thmiddle"),
    ("A_THJ1", "joint", ("THJ1",), "thdistal"),
    ("A_FFJ4", "joint", ("FFJ4",), "knuckle"),
    ("A_FFJ3", "joint", ("FFJ3",), "proximal"),
    ("A_FFJ0", "tendon", ("FFJ2", "FFJ1"), "middle_distal"),
    ("A_MFJ4", "joint", ("MFJ4",), "knuckle"),
    ("A_MFJ3", "joint", ("MFJ3",), "proximal"),
    ("A_MFJ0", "tendon", ("MFJ2", "MFJ1"), "middle_distal"),
    ("A_RFJ4", "joint", ("RFJ4",), "knuckle"),
    ("A_RFJ3", "joint", ("RFJ3",), "proximal"),
    ("A_RFJ0", "tendon", ("RFJ2", "RFJ1"), "middle_distal"),
    ("A_LFJ5", "joint", ("LFJ5",), "metacarpal"),
    ("A_LFJ4", "joint", ("LFJ4",), "knuckle"),
    ("A_LFJ3", "joint", ("LFJ3",), "proximal"),
    ("A_LFJ0", "tendon", ("LFJ2", "LFJ1"), "middle_distal"),
]

# forearm DoFs, reference shadow_hand.py:41-55 (only the default two are supported)
FOREARM_DOFS = {
    "forearm_tx": dict(type="slide", axis=(-1.0, 0.0, 0.0), stiffness=300.0, range=(-1.0, 1.0)),
    "forearm_ty": dict(type="slide", axis=(0.0, 0.0, 1.0), stiffness=300.0, range=(0.0, 0.06)),
}
