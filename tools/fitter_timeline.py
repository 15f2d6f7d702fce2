"""CUPTI timeline of ONE replay of fit.GraphFitter's captured training iteration (projection .. Adan step) at a config:
kernel / memset / memcpy activities with their durations and the gaps between them.   python tools/fitter_timeline.py [C2]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from torch.profiler import ProfilerActivity, profile

from gaussianimage_b200.fit import GraphFitter
from gaussianimage_b200.synth import make_config

cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
s = make_config(cfg)
gt = s.gt.permute(2, 0, 1).unsqueeze(0).contiguous().to("cuda:0")
f = GraphFitter(gt, s.means2d.shape[0])
f.capture()
f.run(20)
torch.cuda.synchronize()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        flush.zero_()
        f.graph.replay()
    torch.cuda.synchronize()
evs = sorted((e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA), key=lambda e: e.time_range.start)
idx = max(i for i, e in enumerate(evs) if "FillFunctor<unsigned char>" in e.name or ("Memset" in e.name and e.time_range.elapsed_us() > 20))
run = evs[idx + 1:]
t0, prev = run[0].time_range.start, run[0].time_range.start
for e in run:
    print(f"+{e.time_range.start - t0:8.1f} us  dur {e.time_range.elapsed_us():7.1f}  gap {e.time_range.start - prev:6.1f}  {e.name[:90]}")
    prev = e.time_range.end
print(f"total {run[-1].time_range.end - t0:.1f} us")
