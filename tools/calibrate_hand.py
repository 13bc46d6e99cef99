"""Replay-based probe of the re-specified Shadow Hand (scene/shadow_hand_model.py).

The reference's recorded policy (examples/twinkle_twinkle_actions.npy) plays TwinkleTwinkleRousseau on the REAL
menagerie hand (tutorial.ipynb cells 15-19). Replaying it on the fp64 oracle with a candidate hand description
and scoring key presses per finger (MidiEvaluationWrapper metrics) says which recollected numbers are off: a
finger that never reaches its key under a policy that was trained to reach it points at that finger's
kinematics. Each hypothesis is a python snippet that edits `shm` (the hand description module) before the
scene is compiled. Test infrastructure: runs the oracle, never the product path.

usage: python tools/calibrate_hand.py hypotheses.py   (a file defining H = {name: snippet}; parallel processes)
A snippet sees `shm`, `np` and `patch_tree(fn)`: fn(tree, find) edits the body tree, find(tree, body_name) -> node."""
import sys
from concurrent.futures import ProcessPoolExecutor
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def run(item):
    name, snippet = item
    import bench
    from diffusion_piano_b200.scene import shadow_hand_model as shm
    from oracle.task_oracle import PianoWithShadowHandsOracle
    sys.path.insert(0, str(ROOT / "tools"))
    from replay_f1 import prf

    def find(node, name):
        if node["name"] == name:
            return node
        for c in node["children"]:
            r = find(c, name)
            if r:
                return r
        return None

    def patch_tree(fn):
        orig = shm.right_hand_tree

        def patched():
            tree = orig()
            fn(tree, find)
            return tree
        shm.right_hand_tree = patched

    exec(snippet, dict(shm=shm, np=np, patch_tree=patch_tree))
    task = bench.build_task()
    acts = bench.recorded_actions()[:158]
    ora = PianoWithShadowHandsOracle(task)
    ora.reset()
    goal, pairs = np.asarray(task.song.goal), np.asarray(task.song.pairs)
    hit, tot, key, ret = np.zeros(10), np.zeros(10), [], 0.0
    off = [[] for _ in range(10)]
    for t in range(158):
        _, r, _ = ora.step(ora.canonical_to_spec(acts[t]))
        ret += r
        act = np.asarray(ora.activation, np.float64)
        key.append(prf(goal[t, :88].astype(np.float64), act))
        for k, f in pairs[t]:
            if k >= 0:
                tot[f] += 1
                hit[f] += act[k]
                off[f].append(ora.physics.site_xpos[f] - ora.physics.geom_xpos[k] - ora.key_geom_pos_offset[k])
    p, r_, f1 = np.mean(key, 0)
    per = " ".join(f"{int(h)}/{int(n)}" for h, n in zip(hit, tot))
    dy = " ".join(f"{1e3 * np.mean(np.array(o)[:, 1]):+.0f}" if o else "-" for o in off)
    dz = " ".join(f"{1e3 * np.mean(np.array(o)[:, 2]):+.0f}" if o else "-" for o in off)
    per += f" | tip-target dy mm: {dy} | dz mm: {dz}"
    return f"{name:28s} P {p:.3f} R {r_:.3f} F1 {f1:.3f} return {ret:7.2f} | per finger (rh th..lf, lh th..lf): {per}"


if __name__ == "__main__":
    ns = {}
    exec(open(sys.argv[1]).read(), ns)
    items = list(ns["H"].items())
    with ProcessPoolExecutor(max_workers=min(8, len(items))) as ex:
        for line in ex.map(run, items):
            print(line, flush=True)
