"""Generates tests/golden/piano_layout.json by RUNNING the reference's own
robopianist/models/piano/piano_mjcf.build() with a recording stub in place of dm_control.mjcf
(dm_control is not installable here; the builder only calls RootElement/add/attribute sets).
Run in the build container only (needs /root/reference)."""
import importlib.util
import json
import sys
import types
from pathlib import Path

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parents[1] / "tests" / "golden" / "piano_layout.json"


class Elem:
    def __init__(self, tag="root", **kw):
        object.__setattr__(self, "_tag", tag)
        object.__setattr__(self, "_attrs", dict(kw))
        object.__setattr__(self, "_children", [])

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        attrs = object.__getattribute__(self, "_attrs")
        if name not in attrs:
            attrs[name] = Elem(name)
        return attrs[name]

    def __setattr__(self, name, value):
        self._attrs[name] = value

    def add(self, tag, **kw):
        e = Elem(tag, **kw)
        self._children.append(e)
        return e


def main():
    mjcf = types.ModuleType("dm_control.mjcf")
    mjcf.RootElement = Elem
    dm = types.ModuleType("dm_control")
    dm.mjcf = mjcf
    mu = types.ModuleType("mujoco_utils")
    mu.types = types.ModuleType("mujoco_utils.types")
    mu.types.MjcfRootElement = object
    sys.modules.update({"dm_control": dm, "dm_control.mjcf": mjcf, "mujoco_utils": mu,
                        "mujoco_utils.types": mu.types})
    for name in ("robopianist", "robopianist.models", "robopianist.models.piano"):
        sys.modules[name] = types.ModuleType(name)
    def load(modname, path):
        spec = importlib.util.spec_from_file_location(modname, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod
    consts = load("robopianist.models.piano.piano_constants", REF / "robopianist/models/piano/piano_constants.py")
    sys.modules["robopianist.models.piano"].piano_constants = consts
    builder = load("robopianist.models.piano.piano_mjcf", REF / "robopianist/models/piano/piano_mjcf.py")
    root = builder.build(add_actuators=True)
    wb = root.worldbody
    bodies = [c for c in wb._children if c._tag == "body"]
    keys = []
    for b in bodies[1:]:
        cls = b._children[0]._attrs["dclass"]
        d = root.default._children[0 if cls == "white_key" else 1]
        keys.append(dict(name=b._attrs["name"], pos=list(map(float, b._attrs["pos"])), dclass=cls,
                         size=list(map(float, d.geom._attrs["size"])), mass=float(d.geom._attrs["mass"]),
                         joint_pos=list(map(float, d.joint._attrs["pos"])),
                         range=list(map(float, d.joint._attrs["range"])),
                         damping=float(d.joint._attrs["damping"]), armature=float(d.joint._attrs["armature"]),
                         stiffness=float(d.joint._attrs["stiffness"]), springref=float(d.joint._attrs["springref"])))
    base = bodies[0]
    out = dict(keys=keys, base=dict(pos=list(map(float, base._attrs["pos"])),
                                    size=list(map(float, base._children[0]._attrs["size"]))),
               n_actuators=len([c for c in root.actuator._children]),
               contype=root.default.geom._attrs["contype"], conaffinity=root.default.geom._attrs["conaffinity"])
    OUT.parent.mkdir(parents=True, exist_ok=True)
    OUT.write_text(json.dumps(out, indent=0))
    print(len(keys), keys[0], keys[1])


if __name__ == "__main__":
    main()
