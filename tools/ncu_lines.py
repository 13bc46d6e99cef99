"""Aggregates an `ncu --page source --csv` SASS dump by CUDA source line, using nvdisasm line info.

usage: python tools/ncu_lines.py <report.ncu-rep> <libpiano_step.so> <kernel substring> [top N] [bucket]
Lines are the OUTERMOST call site (nvdisasm -gi inline chains); bucket > 0 sums over ranges of that many lines.
"""
import csv
import re
import subprocess
import sys
import tempfile
from collections import defaultdict
from pathlib import Path


def main():
    rep, so, kern = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
    bucket = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    tmp = Path(tempfile.mkdtemp())
    subprocess.run(["cuobjdump", "-xelf", "all", str(Path(so).resolve())], cwd=tmp, check=True, capture_output=True)
    lines = None
    for cubin in tmp.glob("*.cubin"):
        sass = subprocess.run(["nvdisasm", "-gi", "-c", str(cubin)], capture_output=True, text=True).stdout.splitlines()
        cur, infunc, out = None, False, []
        for ln in sass:
            if ".section" in ln and ".text." in ln:
                infunc = kern in ln
            m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
            if m:
                cur = (m.group(1), int(m.group(2)))
                continue
            if infunc and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                out.append(cur)
        if out and (lines is None or len(out) > len(lines)):
            lines = out
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))[2:]
    print(f"sass instructions: report {len(rows)}, cubin {len(lines)}")
    agg, samp = defaultdict(float), defaultdict(float)
    for i in range(min(len(rows), len(lines))):
        key = lines[i] or ("?", 0)
        if bucket: key = (key[0], key[1] // bucket * bucket)
        agg[key] += float(rows[i][5]); samp[key] += float(rows[i][4])
    tot, ts = sum(agg.values()), sum(samp.values())
    print(f"warp instructions executed: {tot:.4g}")
    for k, v in sorted(agg.items(), key=lambda x: -x[1])[:top]:
        print(f"{k[0]}:{k[1]:<5d} {100 * v / tot:5.1f}% inst {100 * samp[k] / max(ts, 1):5.1f}% samples")


if __name__ == "__main__":
    main()
