"""Per-kernel counts of the SASS mnemonics that prove what the kernels use (bulk TMA copies, mbarriers, MUFU.EX2, vector
reductions, NVLS multicast stores, warp votes / redux) -> profiles/*_sass_summary.txt.   python tools/sass_summary.py [lib.so]"""
import collections
import re
import subprocess
import sys
from pathlib import Path

lib = sys.argv[1] if len(sys.argv) > 1 else str(Path(__file__).resolve().parent.parent / "gaussianimage_b200" / "libgi2d.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
KEYS = ["UBLKCP", "SYNCS", "MUFU.EX2", "REDG.E.ADD.F32x4", "REDG", "ATOMG", "ATOMS", "VOTE", "REDUX", "CREDUX", "SHFL", "BAR.SYNC", "LDS", "STS", "LDG", "STG", "STG.E.EF", "FFMA", "FMUL"]
cur, counts, total = None, collections.OrderedDict(), {}
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void gi2d::", "")
        counts[cur] = collections.Counter()
        total[cur] = 0
        continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
        ins = line.split("*/", 1)[1].strip()
        ins = re.sub(r"^@!?U?P\d\s+", "", ins)
        total[cur] += 1
        for k in KEYS:
            if ins.startswith(k):
                counts[cur][k] += 1
print(f"# SASS mnemonic counts per kernel of {Path(lib).name} (cuobjdump -sass, sm_100a); static counts, not executed counts")
print("# UBLKCP = cp.async.bulk (1-D bulk TMA copy); SYNCS = mbarrier arrive.expect_tx / try_wait; REDG.E.ADD.F32x4 = red.global.add.v4.f32;")
print("# STG.E.EF = st.global.cs (peer stores of the row-sharded epilogue); the `multimem.st` of the NVLS epilogue is inline PTX that")
print("# cuobjdump prints as a plain STG.E.128 (no distinct mnemonic); CREDUX / REDUX = redux.sync; VOTE = ballots of the mask building.")
print("| kernel | SASS instructions | " + " | ".join(KEYS) + " |")
print("|---|---|" + "---|" * len(KEYS))
for k, c in counts.items():
    print(f"| `{k}` | {total[k]} | " + " | ".join(str(c.get(x, 0)) for x in KEYS) + " |")
