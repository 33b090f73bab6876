"""CPU tests of the multi-GPU host logic (gloo, world_size 2 and 3): slice arithmetic and the variable-length
all-gather that assembles the photon maps / image bands in rank order."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cuda_raytrace_b200 import multigpu as M


def test_path_ranges_partition_exactly():
    for n in (0, 1, 7, 1000, 10_000_019):
        for world in (1, 2, 3, 8):
            parts = [M.path_range(n, r, world) for r in range(world)]
            assert parts[0][0] == 0 and sum(c for _, c in parts) == n
            for (a, ca), (b, _) in zip(parts[:-1], parts[1:]):
                assert a + ca == b


def test_bands_partition_image():
    for yres in (1, 5, 1080):
        for world in (1, 2, 4, 8):
            rows = []
            for r in range(world):
                x0, x1, y0, y1 = M.band_window(yres, 64, r, world)
                assert (x0, x1) == (0, 64)
                rows += list(range(y0, y1))
            assert rows == list(range(yres))


def _worker(rank, world, port, counts, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    local = torch.from_numpy(rng.integers(0, 256, (counts[rank], 32), dtype=np.uint8))
    allp, got_counts = M.allgather_varlen(local)
    q.put((rank, allp.numpy().copy(), got_counts))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("counts", [[5, 9], [0, 4], [3, 0, 7], [0, 0]])
def test_allgather_varlen_rank_order(counts):
    world = len(counts)
    port = 29600 + (os.getpid() + sum(counts) * 7 + world) % 300
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, world, port, counts, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    exp = np.concatenate([np.random.default_rng(100 + r).integers(0, 256, (counts[r], 32), dtype=np.uint8)
                          for r in range(world)]) if sum(counts) else np.zeros((0, 32), np.uint8)
    for rank, allp, got in res:
        assert got == counts
        assert np.array_equal(allp, exp)
