"""Hot spots of one kernel from an ncu report captured with --import-source on: segments between barriers / mbarrier waits
with their share of the stall samples and of the executed instructions, and every instruction above 1.5 % of the samples.
python tools/ncu_source_hot.py gpurun_out/r01_fwd_c5.ncu-rep [kernel-name-regex]"""
import csv
import io
import subprocess
import sys

cmd = ["ncu", "-i", sys.argv[1], "--page", "source", "--csv"]
if len(sys.argv) > 2:          # a report with several kernels: pick one by (regex) name; the first matching launch is shown
    cmd += ["--kernel-name", "regex:" + sys.argv[2], "--launch-count", "1"] if False else ["--kernel-name", "regex:" + sys.argv[2]]
raw = subprocess.run(cmd, capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
data = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    try:
        data.append((int(r[idx["Warp Stall Sampling (All Samples)"]] or 0), int(r[idx["Instructions Executed"]] or 0), r[idx["Source"]].strip()))
    except ValueError:
        pass
tot_s, tot_i = sum(d[0] for d in data), sum(d[1] for d in data)
print(rows[0][1][:100])
print(f"samples {tot_s}, warp instructions {tot_i / 1e6:.1f} M, {len(data)} SASS instructions")
start = cur_s = cur_i = 0
for n, (s, i, src) in enumerate(data):
    cur_s += s
    cur_i += i
    if "BAR.SYNC" in src or "TRYWAIT" in src or n == len(data) - 1:
        print(f"[{start:4d}-{n:4d}] samples {100 * cur_s / tot_s:5.1f} %  instructions {cur_i / 1e6:7.1f} M   ends with {src[:44]}")
        start, cur_s, cur_i = n + 1, 0, 0
for n, (s, i, src) in enumerate(data):
    if s >= tot_s * 0.015:
        print(f"{n:4d} {100 * s / tot_s:5.1f} %  executed {i / 1e6:7.2f} M  {src[:80]}")
