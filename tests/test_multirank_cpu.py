"""Host-side logic of the N > 1 path on CPU: world_size-2 gloo. The band partition every rank derives must tile the
image exactly once, and the rendezvous bench.py uses to ship the 128-byte communicator id (broadcast_object_list from
rank 0) must deliver identical bytes to every rank. (The device side of the same path is tools/mgpu_check.py and
tests/test_gpu_multi.py.)"""
import os
import sys

import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from gs2d_loader import ROOT, gs2d


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    box = [bytes(range(128)) if rank == 0 else None]  # stands in for gs2d_comm_unique_id() (needs NCCL + a GPU)
    dist.broadcast_object_list(box, src=0)
    bands = {}
    for tiles_y in (1, 2, 17, 256, 512):
        bands[tiles_y] = gs2d.band_rows(tiles_y, rank, world)
    gathered = [None] * world
    dist.all_gather_object(gathered, (box[0], bands))
    if rank == 0:
        q.put(gathered)
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_bands_partition_and_id_broadcast(world):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29600 + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    gathered = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ids = {g[0] for g in gathered}
    assert ids == {bytes(range(128))}
    for tiles_y in (1, 2, 17, 256, 512):
        rows = []
        for g in gathered:
            a, b = g[1][tiles_y]
            assert 0 <= a <= b <= tiles_y
            rows += list(range(a, b))
        assert rows == list(range(tiles_y)), (tiles_y, rows)  # every tile row exactly once, in rank order
