"""Probe torch symmetric memory on this box: torchrun --nproc-per-node 2 tools/symm_probe.py"""
import os
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dev = torch.device("cuda", rank)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty((1024, 1024), dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
print(rank, "world", hdl.world_size, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "multicast_ptr", hex(hdl.multicast_ptr or 0), "buffer_size", hdl.buffer_size, flush=True)
t.fill_(rank + 1)
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (1024, 1024), torch.float32)
print(rank, "peer value", float(peer[0, 0]), flush=True)
hdl.barrier(channel=1)
dist.destroy_process_group()
