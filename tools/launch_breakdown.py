"""Per-kernel time of the last control step in an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_breakdown.py <launches.csv> [launches of the last step, default 50]"""
import collections
import csv
import re
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith("==")]
rows = [(x["Kernel Name"], float(x["Metric Value"].replace(",", ""))) for x in csv.DictReader(lines)
        if "gpu__time_duration" in x.get("Metric Name", "")]
hk = [(k, v) for k, v in rows if "pp_hands_step_kernel" in k]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 50
NAMES = {0: "whole step", 1: "prologue", 2: "newton", 3: "finish"}


def mode(k):
    m = re.search(r"pp_hands_step_kernel<\D*(\d+)\D+(\d+)\D*>", k)
    return (NAMES[int(m.group(2))], int(m.group(1)))


last = [(mode(k), v / 1e3) for k, v in hk][-n:]
tot, cnt = collections.Counter(), collections.Counter()
for m, v in last:
    tot[m] += v
    cnt[m] += 1
for m in tot:
    print(f"{m[0]:10s} {m[1]:2d} envs/block: {cnt[m]:3d} launches {tot[m] / 1e3:8.3f} ms total {tot[m] / cnt[m]:8.1f} us avg")
print(f"sum {sum(tot.values()) / 1e3:.3f} ms")
