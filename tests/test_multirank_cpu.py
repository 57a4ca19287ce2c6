"""Host-side multi-rank plumbing on CPU (gloo, world_size 2): the partition plan and the
all-gather wrapper the partitioned solver builds its reduced system with."""
import os
import socket

import numpy as np
import pytest
import torch

from fvm_b200.partitioned import Comm, plan_partition
from fvm_b200 import workloads as W


def test_plan_partition_covers_every_cell_once():
    for J, T, world in [(2 ** 20, 8192, 8), (1000, 256, 3), (8192 * 5 + 1, 8192, 4), (3, 256, 2)]:
        pl = plan_partition(J, T, world)
        assert pl["tiles_global"] == pl["tiles_per_rank"] * world >= -(-J // T)
        r = pl["ranges"]
        assert r[0][0] == 0 and r[-1][1] == J
        assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
        assert all(a % T == 0 or a == J for a, _ in r)


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    comm = Comm(dist.group.WORLD)
    assert (comm.rank, comm.world) == (rank, world)
    t = torch.arange(6, dtype=torch.float64).reshape(2, 3) + 10 * rank
    g = comm.all_gather(t)
    assert g.shape == (world, 2, 3)
    for r in range(world):
        assert torch.equal(g[r], torch.arange(6, dtype=torch.float64).reshape(2, 3) + 10 * r)
    assert comm.all_reduce_max_int([rank, 5 - rank]) == [world - 1, 5]
    # shard-invariant workload: each rank generates its own shard, together they are the full batch
    wl = W.batched_sweep(8, 16, first=4 * rank, count=4)
    allk = comm.all_gather(torch.as_tensor(wl["k"]))
    np.save(os.path.join(out_dir, "k%d.npy" % rank), allk.reshape(-1).numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_comm_all_gather_gloo_world2(tmp_path):
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    full = W.batched_sweep(8, 16)["k"]
    for r in range(2):
        assert np.array_equal(np.load(tmp_path / ("k%d.npy" % r)), full)


def test_blocks_of_the_separator_system():
    """How the host layer cuts a rank's part of the separator system into blocks (partitioned.pick_blocks)."""
    from fvm_b200.partitioned import pick_blocks
    assert pick_blocks(32768, 1) == 128           # cfg 5 on one GPU: 128 blocks of 256 separators
    assert pick_blocks(32768, 8) == 128           # ... and on eight: 1,024 blocks in all, the top system's limit
    assert pick_blocks(32768, 16) == 64
    assert pick_blocks(256, 1) == 1 and pick_blocks(511, 1) == 1 and pick_blocks(512, 1) == 2
    assert pick_blocks(3 * 257, 1) == 3           # only divisors of the tile count
    assert pick_blocks(12, 1, wanted=3) == 3 and pick_blocks(4, 2, wanted=2) == 2
    for nt, world, wanted in ((12, 1, 5), (12, 1, 12), (12, 1, 0), (4096, 8, 256)):
        with pytest.raises(ValueError):
            pick_blocks(nt, world, wanted)
