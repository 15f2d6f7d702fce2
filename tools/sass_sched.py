"""Prints a kernel's SASS with the scheduling control bits decoded (stall count, yield, barriers) -- to see how well the
hot loops are interleaved.   python tools/sass_sched.py <kernel-name-regex> [lib.so] > out.txt"""
import re
import subprocess
import sys
from pathlib import Path

pat = re.compile(sys.argv[1])
lib = sys.argv[2] if len(sys.argv) > 2 else str(Path(__file__).resolve().parent.parent / "gaussianimage_b200" / "libgi2d.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout.splitlines()
on = False
i = 0
while i < len(sass):
    line = sass[i]
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        on = bool(pat.search(name))
        if on:
            print("==", name[:160])
    elif on:
        m = re.match(r"\s+/\*([0-9a-f]{4})\*/\s+(.*?)\s*/\* 0x([0-9a-f]{16}) \*/", line)
        if m and i + 1 < len(sass):
            m2 = re.search(r"/\* 0x([0-9a-f]{16}) \*/", sass[i + 1])
            hi = int(m2.group(1), 16)
            ctrl = hi >> 41
            stall, yld, wr, rd, wait = ctrl & 15, (ctrl >> 4) & 1, (ctrl >> 5) & 7, (ctrl >> 8) & 7, (ctrl >> 11) & 63
            print(f"{m.group(1)}  st{stall:2d} {'Y' if yld else ' '} w{wr if wr != 7 else '-'} r{rd if rd != 7 else '-'} wt{wait:02x}  {m.group(2)}")
            i += 1
        elif re.match(r"\s+\.L_", line):
            print(line.strip())
    i += 1
