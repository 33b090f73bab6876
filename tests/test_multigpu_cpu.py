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


@pytest.mark.parametrize("yres,world,stripes", [(1080, 8, None), (2160, 2, None), (2160, 8, 4), (17, 4, 8), (5, 8, None), (64, 1, None),
                                                (1000, 3, 7)])
def test_interleaved_stripes_partition_and_compose(yres, world, stripes):
    """render_sharded's band plan: the bands tile the rows exactly, band b belongs to rank b % world, and stacking each
    rank's bands then composing them gives the image back (numpy stand-in for the gathered CUDA tensors)."""
    edges = M.stripe_plan(yres, world, stripes)
    assert edges[0] == 0 and edges[-1] == yres and all(a <= b for a, b in zip(edges, edges[1:]))
    assert (len(edges) - 1) % world == 0
    img = np.arange(yres * 3, dtype=np.float32).reshape(yres, 3)
    covered = np.zeros(yres, np.int32)
    parts = []
    for r in range(world):
        rows = M.stripe_rows(edges, world, r)
        for y0, y1 in rows:
            covered[y0:y1] += 1
        parts.append(np.concatenate([img[y0:y1] for y0, y1 in rows] + [np.zeros((0, 3), np.float32)], 0))
    assert (covered == 1).all()
    assert np.array_equal(np.concatenate(M.stripe_compose(parts, edges, world), 0), img)
    if stripes is None and yres >= 64 * world:
        assert min(b - a for a, b in zip(edges, edges[1:])) >= 64 - 1          # default: bands of (about) 64 rows or more


def _fg_worker(rank, world, port, n_paths, q):
    """One rank of the sharded second photon pass of final gathering, with the C oracle standing in for the GPU tracer:
    trace this rank's slice of the path range, assemble both maps in rank order."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import oracle as O
    from cuda_raytrace_b200 import scenes as S
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    first, count = M.path_range(n_paths, rank, world)
    d, r = sc.photon_trace_fg(7, first, count)
    out = []
    for part in (d, r):
        local = torch.from_numpy(np.ascontiguousarray(part).view(np.uint8).reshape(-1, 32).copy())
        allp, counts = M.allgather_varlen(local)
        out.append((allp.numpy().copy(), counts))
    q.put((rank, out))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_final_gather_photon_pass_equals_single_rank(world):
    """multigpu.shoot_sharded_fg's contract on CPU: direct and radiance photons traced by path-range slices and
    concatenated in rank order are the single-rank arrays byte for byte (a path's photons depend only on seed and index)."""
    from oracle import oracle as O
    from cuda_raytrace_b200 import scenes as S
    n_paths = 3001
    port = 29300 + (os.getpid() + world * 13) % 250
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_fg_worker, args=(r, world, port, n_paths, q)) for r in range(world)]
    for p in ps:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    d, r = sc.photon_trace_fg(7, 0, n_paths)
    assert len(d) > 1000 and len(r) > 200
    for rank, out in res:
        for (allp, counts), ref in zip(out, (d, r)):
            assert sum(counts) == len(ref) and len(counts) == world
            assert np.array_equal(allp, np.ascontiguousarray(ref).view(np.uint8).reshape(-1, 32))
