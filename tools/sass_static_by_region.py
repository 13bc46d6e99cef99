"""Static SASS instruction count per source region of a kernel (nvdisasm --print-line-info).
Usage: python tools/sass_static_by_region.py <cubin> <kernel substring> [region size in lines]
Shows where the code BYTES are (16 B per instruction) - what the instruction caches have to hold."""
import re
import subprocess
import sys
from collections import Counter

cubin, kern = sys.argv[1], sys.argv[2]
step = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout.splitlines()
inside, cur, cnt, total = False, ("?", 0), Counter(), 0
for ln in txt:
    if ln.startswith(".text."):
        inside = kern in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", ln):
        cnt[(cur[0], cur[1] // step * step)] += 1
        total += 1
print(f"kernel ~{kern}: {total} instructions = {total * 16 / 1024:.0f} KB")
for (f, l), n in sorted(cnt.items()):
    print(f"{f}:{l:5d}  {n:6d}  {n * 16 / 1024:6.1f} KB")
