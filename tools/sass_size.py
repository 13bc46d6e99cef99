"""SASS instruction count of every function in libpiano_step.so, and of the hands kernel by source region.

usage: python tools/sass_size.py [bucket]
"""
import re
import subprocess
import sys
import tempfile
from collections import Counter
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def main():
    bucket = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    tmp = Path(tempfile.mkdtemp())
    subprocess.run(["cuobjdump", "-xelf", "all", str(ROOT / "diffusion-piano_b200" / "libpiano_step.so")], cwd=tmp,
                   check=True, capture_output=True)
    for cubin in sorted(tmp.glob("*.cubin")):
        sass = subprocess.run(["nvdisasm", "-g", "-c", str(cubin)], capture_output=True, text=True).stdout.splitlines()
        func, cur = None, None
        per_func, per_region = Counter(), Counter()
        for ln in sass:
            if ".section" in ln and ".text." in ln:
                func = ln.split(".text.")[1].split(",")[0][:60]
            m = re.search(r'//## File ".*?([^/"]+)", line (\d+)', ln)
            if m:
                cur = (m.group(1), int(m.group(2)))
                continue
            if func and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln):
                per_func[func] += 1
                if bucket and "pp_hands_step" in func and cur:
                    key = (cur[0], cur[1] // bucket * bucket if cur[0] == "pp_hands.cu" else 0)
                    per_region[key] += 1
        for f, n in per_func.most_common():
            print(f"{cubin.name:40s} {f:62s} {n}")
        for k, n in per_region.most_common(30):
            print("   ", k, n)


if __name__ == "__main__":
    main()
