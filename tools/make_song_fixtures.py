"""Regenerates diffusion-piano_b200/data/*.json|npz from the reference checkout.

Run in the build container only (needs /root/reference). The outputs are data, not code:
tick-level MIDI events re-encoded as JSON (so the package's own SMF->NoteSequence rules
run on them at load time), the hard-coded fingering lists of the two Rousseau songs
(reference robopianist/music/library.py:350-385 and the nocturne list), the fingering
annotation rows, and the recorded action sequence examples/twinkle_twinkle_actions.npy.
"""
import ast
import importlib
import json
import re
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference")
smf = importlib.import_module("diffusion_piano_b200.music.smf")
OUT = ROOT / "diffusion-piano_b200" / "data"


def events(path):
    res, tracks = smf.parse_smf_events(Path(path).read_bytes())
    keep = ("set_tempo", "time_signature", "note_on", "note_off", "control_change", "program_change")
    return {"resolution": res, "tracks": [[list(e) for e in t if e[1] in keep] for t in tracks]}


def fingering_lists():
    src = (REF / "robopianist/music/library.py").read_text()
    out = []
    for m in re.finditer(r"FINGERING = \[(.*?)\]", src, re.S):
        body = re.sub(r"#.*", "", m.group(1))
        out.append(list(ast.literal_eval("[" + body + "]")))
    return out


def annotation_rows(path):
    rows = []
    for line in Path(path).read_text().splitlines():
        if line.startswith("//") or not line.strip():
            continue
        rows.append(line.strip().split("\t"))
    return rows


def main():
    OUT.mkdir(exist_ok=True, parents=True)
    fl = fingering_lists()
    songs = {
        "twinkle_twinkle_rousseau": dict(
            events(REF / "robopianist/music/data/rousseau/twinkle-twinkle-trimmed.mid"),
            fingering=fl[0], artist="Rousseau", title="Twinkle Twinkle (YouTube)"),
        "nocturne_rousseau": dict(
            events(REF / "robopianist/music/data/rousseau/nocturne-trimmed.mid"),
            fingering=fl[1], artist="Rousseau", title="Nocturne (YouTube)"),
        "guren_no_yumiya_cut_14s": dict(
            events(REF / "midi_files_cut/Guren no Yumiya Cut 14s.mid"),
            annotation=annotation_rows(REF / "data_processing/Guren no Yumiya Cut 14s_fingering v3.txt")),
        "crossing_field_cut_10s": dict(
            events(REF / "midi_files_cut/Crossing Field Cut 10s.mid"),
            annotation=annotation_rows(REF / "data_processing/Crossing_Field_Fingering.txt")),
    }
    (OUT / "songs.json").write_text(json.dumps(songs, separators=(",", ":")))
    a = np.load(REF / "examples/twinkle_twinkle_actions.npy")
    np.savez_compressed(OUT / "recorded_actions.npz", twinkle_twinkle=a)
    print({k: sum(len(t) for t in v["tracks"]) for k, v in songs.items()}, a.shape)


if __name__ == "__main__":
    main()
