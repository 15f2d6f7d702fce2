"""Stage timing of the row-sharded decode (CUDA events, per rank).  torchrun --nproc-per-node N tools/decode_breakdown.py [fused]"""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from gaussianimage_b200 import cuda as GC
from gaussianimage_b200.sharded import RowShardedRenderer
from gaussianimage_b200.synth import make_config

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
fused = len(sys.argv) > 1 and sys.argv[1] == "fused"
s = make_config("C5")
d = s.to(dev)
R = RowShardedRenderer(s.H, s.W, dev, fused=fused and world > 1)
N = d.means2d.shape[0]


def ev():
    return torch.cuda.Event(enable_timing=True)


acc = {}
for it in range(8):
    e = [ev() for _ in range(5)]
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e[0].record()
    xys, depths, radii, conics, nth = GC.project_gaussians_2d_forward(N, d.means2d, d.cholesky, s.H, s.W, (R.TX, R.TY, 1))
    e[1].record()
    if R.fused:
        R._hdl.barrier(channel=0)
    e[2].record()
    fz, R.fused = R.fused, False           # render_band without its own barrier
    R.render_band(xys, radii, conics, nth, d.colors, d.opacity)
    R.fused = fz
    e[3].record()
    R.gather()
    e[4].record()
    torch.cuda.synchronize()
    if it >= 3:
        for k, (a, b) in {"project": (0, 1), "barrier0": (1, 2), "bin+raster": (2, 3), "gather/barrier1": (3, 4), "total": (0, 4)}.items():
            acc.setdefault(k, []).append(e[a].elapsed_time(e[b]) * 1e3)
print(f"rank {rank}/{world} fused={R.fused}: " + ", ".join(f"{k} {sum(v) / len(v):.0f} us" for k, v in acc.items()), flush=True)
if world > 1:
    dist.destroy_process_group()
