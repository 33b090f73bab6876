"""BASELINE configs[4]: photon-count sweep on the Cfg3 scene.  For each emitted-path count: emit + trace
(sharded by path range over the ranks), photon-map assembly (all-gather), grid build (replicated), camera pass with
the kNN gather at 1920x1080, 1 spp (bands over the ranks).  Stage times are wall clock around synchronised stages, max
over ranks, best of `reps` after one warm-up.  Prints one JSON object per sweep point on rank 0.
  python tools/bench_cfg5.py [paths ...]            (default 2e6 2e7 2e8; ~0.52 photons are stored per path)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_cfg5.py ..."""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S, multigpu as M   # noqa: E402

sweep = [int(float(a)) for a in sys.argv[1:]] or [2_000_000, 20_000_000, 200_000_000]
rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
BATCH = 24_000_000      # paths per b2rt_photons_trace call (slot buffer = 5 x 32 B per path)

ctx = b2rt.Context(local, seed=3)
S.fill_scene(ctx, S.cornell_scene(spheres=True, wide=1.78))
ctx.build_accel()
cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 1920, 1080)

frame_out = ctx.pinned_empty((1080, 1920, 3))

def sync():
    ctx.sync()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def shoot(n_paths):
    """this rank's slice of [0, n_paths) in batches; returns nothing (photons stay in the context)"""
    first, count = M.path_range(n_paths, rank, world)
    ctx.photons_clear()
    for f in range(first, first + count, BATCH):
        ctx.photons_trace(f, min(BATCH, first + count - f))


def assemble(n_paths):
    """all-gather of both maps (the second half of multigpu.shoot_sharded)"""
    return [M.assemble_map(ctx, which, n_paths) for which in (b2rt.MAP_CAUSTIC, b2rt.MAP_INDIRECT)]


for n_paths in sweep:
    reps = 3 if n_paths <= 50_000_000 else 2
    best = None
    for rep in range(reps + 1):
        sync(); t0 = time.perf_counter()
        shoot(n_paths)
        sync(); t1 = time.perf_counter()
        if world > 1:
            nc, ni = assemble(n_paths)
        else:
            nc, ni = ctx.photons_count(0)[0], ctx.photons_count(1)[0]
        sync(); t2 = time.perf_counter()
        g0 = ctx.grid_build(0); g1 = ctx.grid_build(1)
        sync(); t3 = time.perf_counter()
        if world > 1:
            M.render_sharded(ctx, cam, spp=1, nused=64, max_dist=0.1)
        else:
            ctx.render_frame(cam, spp=1, nused=64, max_dist=0.1, out=frame_out)
        sync(); t4 = time.perf_counter()
        tt = torch.tensor([t1 - t0, t2 - t1, t3 - t2, t4 - t3, t4 - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        tt = (tt * 1e3).tolist()
        if rep >= 1 and (best is None or tt[4] < best[4]):
            best = tt
    if rank == 0:
        n = nc + ni
        gathered = 32.0 * n * (world - 1) / world        # bytes each rank receives
        print(json.dumps({"n_gpus": world, "paths": n_paths, "photons": n, "caustic": nc, "indirect": ni,
                          "shoot_ms": best[0], "allgather_ms": best[1], "grid_ms": best[2], "camera_gather_ms": best[3],
                          "frame_ms": best[4], "paths_per_s": n_paths / (best[0] * 1e-3),
                          "photons_traced_and_gathered_per_s": n / (best[4] * 1e-3),
                          "allgather_bus_GBs": (gathered / (best[1] * 1e-3) / 1e9) if world > 1 else None,
                          "grid_cells": [list(g0.dims), list(g1.dims)], "cell_size": [g0.cell_size, g1.cell_size]}), flush=True)
ctx.close()
if world > 1:
    dist.destroy_process_group()
