"""Coordinate search over uncertain Shadow Hand numbers, scored by the replay of the reference's recorded policy
(tools/calibrate_hand.py). Prints every evaluation; the best point of each sweep seeds the next.
usage: python tools/calibrate_search.py [sweeps]"""
import math
import re
import sys
from concurrent.futures import ProcessPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
import calibrate_hand as ch  # noqa: E402

SPACE = {
    "lfj5_phi": [35, 60, 90, 105, 120, 135],          # LFJ5 axis = (sin phi, 0, cos phi)
    "rflf_j4_flip": [0, 1],                           # RFJ4 / LFJ4 axis (0, 1, 0) instead of the class (0, -1, 0)
    "tip_r": [0.0065, 0.0075, 0.0085, 0.0095],        # finger distal capsule radius
    "tip_hl": [0.006, 0.008, 0.010],
    "tip_z": [0.012, 0.0135, 0.015],
    "th_r": [0.009, 0.011, 0.013],                    # thumb distal capsule
    "th_hl": [0.007, 0.009, 0.011],
    "th_z": [0.014, 0.016, 0.018],
    "thbase_deg": [30, 45, 60],                       # thbase tilt about y
    "th_kp_scale": [1.0, 2.0],
    "site_dz": [0.0],
}
BASE = dict(lfj5_phi=35, rflf_j4_flip=0, tip_r=0.0075, tip_hl=0.008, tip_z=0.0135, th_r=0.009, th_hl=0.009,
            th_z=0.016, thbase_deg=45, th_kp_scale=1.0, site_dz=0.0)


def snippet(p):
    a = (round(math.sin(math.radians(p["lfj5_phi"])), 6), 0.0, round(math.cos(math.radians(p["lfj5_phi"])), 6))
    h = math.radians(p["thbase_deg"]) / 2
    lines = ["def _e(t, f):",
             f"    f(t, 'lfmetacarpal')['joint']['axis'] = {a}",
             f"    f(t, 'thbase')['quat'] = ({math.cos(h):.6f}, 0.0, {math.sin(h):.6f}, 0.0)",
             f"    f(t, 'thdistal')['geoms'][0] = dict(type='capsule', size=({p['th_r']}, {p['th_hl']}), pos=(0, 0, {p['th_z']}))",
             "    for n in ('ffdistal', 'mfdistal', 'rfdistal', 'lfdistal'):",
             f"        f(t, n)['geoms'][0] = dict(type='capsule', size=({p['tip_r']}, {p['tip_hl']}), pos=(0, 0, {p['tip_z']}))"]
    if p["rflf_j4_flip"]:
        lines += ["    f(t, 'rfknuckle')['joint']['axis'] = (0.0, 1.0, 0.0)",
                  "    f(t, 'lfknuckle')['joint']['axis'] = (0.0, 1.0, 0.0)"]
    lines += ["patch_tree(_e)"]
    if p["th_kp_scale"] != 1.0:
        lines += [f"for c in ('thbase', 'thproximal', 'thhub', 'thmiddle', 'thdistal'): shm.ACT_CLASS[c] = dict(shm.ACT_CLASS[c], kp=shm.ACT_CLASS[c]['kp'] * {p['th_kp_scale']})"]
    return "\n".join(lines) + "\n"


def score(line):
    m = re.search(r"F1 ([0-9.]+) return\s+([0-9.]+)", line)
    return float(m.group(1)) + 1e-4 * float(m.group(2))


def main():
    sweeps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
    best = dict(BASE)
    with ProcessPoolExecutor(max_workers=8) as ex:
        best_line = list(ex.map(ch.run, [("start", snippet(best))]))[0]
        print(best_line, flush=True)
        for s in range(sweeps):
            for name, values in SPACE.items():
                cands = [dict(best, **{name: v}) for v in values if v != best[name]]
                if not cands:
                    continue
                lines = list(ex.map(ch.run, [(f"{name}={c[name]}", snippet(c)) for c in cands]))
                for c, ln in zip(cands, lines):
                    print(ln.split("| tip-target")[0], flush=True)
                    if score(ln) > score(best_line):
                        best, best_line = c, ln
                print(f"== sweep {s} after {name}: best {best}  F1-score {score(best_line):.4f}", flush=True)
    print("BEST", best)
    print(best_line)


if __name__ == "__main__":
    main()
