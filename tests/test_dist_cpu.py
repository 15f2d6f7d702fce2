"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: tile-row bands + in-place all-gather assemble the same
frame as a single process; image-parallel sharding needs no collective."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gaussianimage_b200.synth import make_scene
from oracle import gi2d_oracle as O


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, H, W, N, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from gaussianimage_b200.sharded import RowShardedRenderer, image_shard
    s = make_scene(H, W, N, seed=1)
    full = O.render_cholesky(*[t.numpy() for t in (s.means2d, s.cholesky, s.colors, s.opacity)], H, W)
    R = RowShardedRenderer(H, W, "cpu")
    assert R.world == world and R.rank == rank
    r0, r1 = R.bands[rank]
    # this rank "renders" only its band (rows of the oracle frame stand in for the CUDA kernels on a CPU box)
    band = torch.from_numpy(full[r0 * 16:min(r1 * 16, H)])
    R.buffer[rank * R.band_rows: rank * R.band_rows + band.shape[0]] = band
    frame = R.gather()
    ok = bool(np.array_equal(frame.numpy(), full))
    mine = image_shard(24, rank, world)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    q.put((rank, ok, sorted(i for g in gathered for i in g) == list(range(24))))
    dist.destroy_process_group()


def test_row_sharded_gather_world2_matches_single_process():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 70, 100, 300, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]       # generous: the driver may run this beside a build on a loaded box
    for p in procs:
        p.join(timeout=60)
    assert all(p.exitcode == 0 for p in procs)
    assert all(ok and shard_ok for _, ok, shard_ok in res), res
