"""Per-loop accounting from an `ncu --page source --csv` export of ONE launch: finds backward branches, reports for every loop
its executed warp instructions, average active threads and pc samples; and the instruction count outside any loop.
  python tools/ncu_loops.py <src.csv> [min_share_percent]"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ia, isrc, ismp, iex, ithr = h.index("Address"), h.index("Source"), h.index("# Samples"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
ins = []
for r in rows[hi + 1:]:
    if r and r[0] in ("Address", "Kernel Name", "#"):      # the export continues with the CUDA-C view of the same launch: SASS only
        break
    if len(r) <= iex or not r[ia].startswith("0x"):
        continue
    ins.append((int(r[ia], 16), r[isrc].strip(), int(r[ismp] or 0), int(r[iex] or 0), float(r[ithr] or 0)))
base = ins[0][0]
addr_idx = {a: k for k, (a, *_rest) in enumerate(ins)}
total = sum(x[3] for x in ins)
tot_s = sum(x[2] for x in ins)
loops = []
for k, (a, s, smp, ex, thr) in enumerate(ins):
    m = re.search(r"BRA(?:\.\w+)*\s+(?:\S+,\s*)?(0x[0-9a-f]+)", s)
    if m:
        t = int(m.group(1), 16)
        if t in addr_idx and addr_idx[t] <= k and ex > 0:
            loops.append((addr_idx[t], k))
print(f"{len(ins)} SASS instructions, {total / 1e6:.2f} M executed warp instructions, {tot_s} samples")
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
for lo, hi_ in sorted(set(loops)):
    ex = sum(x[3] for x in ins[lo:hi_ + 1])
    smp = sum(x[2] for x in ins[lo:hi_ + 1])
    if 100.0 * ex / total < min_share:
        continue
    thr = sum(x[3] * x[4] for x in ins[lo:hi_ + 1]) / max(ex, 1)
    trips = ins[hi_][3]
    mufu = sum(x[3] for x in ins[lo:hi_ + 1] if "MUFU.EX2" in x[1])
    print(f"loop [{lo:5d}-{hi_:5d}] {hi_ - lo + 1:5d} instr  executed {ex / 1e6:6.2f} M = {100.0 * ex / total:5.1f} %  samples {100.0 * smp / max(tot_s, 1):5.1f} %  "
          f"avg active threads {thr:4.1f}  back-branch executed {trips}  MUFU.EX2 {mufu}")
