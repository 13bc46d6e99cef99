"""Warp-stall samples of an `ncu --set full --import-source on` report by source region and stall reason.

usage: python tools/ncu_stalls.py <report.ncu-rep> <libpiano_step.so> <kernel substring> [bucket lines] [top N] [max line]
Lines are the OUTERMOST call site of inlined code (nvdisasm -gi), like tools/ncu_lines.py; with [max line], the outermost
one BELOW that line (the step body is itself inlined into the kernel wrappers near the end of pp_hands.cu).
"""
import csv
import re
import subprocess
import sys
import tempfile
from collections import defaultdict
from pathlib import Path

REASONS = ["stall_wait", "stall_short_sb", "stall_no_inst", "stall_branch_resolving", "stall_long_sb",
           "stall_not_selected", "stall_selected", "stall_math", "stall_barrier", "stall_dispatch"]


def main():
    rep, so, kern = sys.argv[1:4]
    bucket = int(sys.argv[4]) if len(sys.argv) > 4 else 20
    top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    max_line = int(sys.argv[6]) if len(sys.argv) > 6 else 1 << 30
    tmp = Path(tempfile.mkdtemp())
    subprocess.run(["cuobjdump", "-xelf", "all", str(Path(so).resolve())], cwd=tmp, check=True, capture_output=True)
    lines = None
    for cubin in tmp.glob("*.cubin"):
        sass = subprocess.run(["nvdisasm", "-gi", "-c", str(cubin)], capture_output=True, text=True).stdout.splitlines()
        cur, infunc, out, fresh = None, False, [], True
        for ln in sass:
            if ".section" in ln and ".text." in ln:
                infunc = kern in ln
            m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
            if m:
                if int(m.group(2)) < max_line or cur is None or fresh:
                    cur = (m.group(1), int(m.group(2)))
                fresh = False
                continue
            if re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                fresh = True  # the next //## File line starts a new inline chain
                if infunc:
                    out.append(cur)
        if out and (lines is None or len(out) > len(lines)):
            lines = out
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, rows = rows[1], rows[2:]
    col = {name: hdr.index(name) for name in REASONS + ["# Samples", "Instructions Executed"]}
    print(f"sass instructions: report {len(rows)}, cubin {len(lines)}")
    agg = defaultdict(lambda: defaultdict(float))
    for i in range(min(len(rows), len(lines))):
        f, l = lines[i] or ("?", 0)
        key = (f, l // bucket * bucket)
        for name, c in col.items():
            try:
                agg[key][name] += float(rows[i][c] or 0)
            except ValueError:
                pass
    tot = defaultdict(float)
    for v in agg.values():
        for k, x in v.items():
            tot[k] += x
    ts = max(tot["# Samples"], 1.0)
    short = [r.replace("stall_", "")[:10] for r in REASONS]
    print(f"{'region':24s} {'%smp':>6s} {'%inst':>6s} " + " ".join(f"{s:>10s}" for s in short) + "   (% of all samples)")
    print(f"{'TOTAL':24s} {100.0:6.1f} {100.0:6.1f} " + " ".join(f"{100 * tot[r] / ts:10.1f}" for r in REASONS))
    for key, v in sorted(agg.items(), key=lambda kv: -kv[1]["# Samples"])[:top]:
        print(f"{key[0] + ':' + str(key[1]):24s} {100 * v['# Samples'] / ts:6.1f} "
              f"{100 * v['Instructions Executed'] / max(tot['Instructions Executed'], 1):6.1f} "
              + " ".join(f"{100 * v[r] / ts:10.2f}" for r in REASONS))


if __name__ == "__main__":
    main()
