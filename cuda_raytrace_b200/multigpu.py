"""Sharding of the path across the GPUs of one box: one process per GPU, torch.distributed for the plumbing.

SURVEY.md 8(e).  The reference has no multi-GPU path at all (one implicit device, cuda_render/cudarender.cpp:17-36).
Two decompositions, each with exactly one exchange step:

  * photon pass  -- rank r traces the path range [r*n/G, (r+1)*n/G) (a path's photons depend only on (seed, path
    index)); the per-rank photon arrays, already in path order, are concatenated in rank order with an all-gather
    (counts first, then the padded payload), so every rank holds the byte-identical G = 1 array and builds the same
    grid.  NCCL over NVLink on the GPU box, gloo on CPU in the tests.
  * camera pass  -- the image is cut into horizontal bands (crop windows of b2rt_render_frame); each pixel is produced
    by exactly one rank with tiling-independent randomness; bands are gathered on rank 0.

No collective is used for plain ray batches (bench.py: every rank traces its own samples).
"""
import numpy as np


def path_range(n_paths, rank, world):
    """[first, first+count) of rank's slice; slices are contiguous, ordered by rank, and cover [0, n_paths)."""
    lo = n_paths * rank // world
    hi = n_paths * (rank + 1) // world
    return lo, hi - lo


def band_window(yres, xres, rank, world):
    """crop window (x0, x1, y0, y1) of rank's horizontal band; empty bands are possible when world > yres"""
    y0 = yres * rank // world
    y1 = yres * (rank + 1) // world
    return 0, xres, y0, y1


def allgather_varlen(local, group=None):
    """local: 2-D torch tensor [n_local, width] (uint8 rows of a record array).  Returns (concatenation over ranks in
    rank order, per-rank counts).  Works with NCCL (CUDA tensors) and gloo (CPU tensors)."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n_local = torch.tensor([local.shape[0]], dtype=torch.int64, device=local.device)
    counts = [torch.zeros_like(n_local) for _ in range(world)]
    dist.all_gather(counts, n_local, group=group)
    counts = [int(c.item()) for c in counts]
    pad = max(max(counts), 1)
    buf = torch.zeros((pad, local.shape[1]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    out = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(out, buf, group=group)
    return torch.cat([o[:c] for o, c in zip(out, counts)], 0), counts


def _ctx_stream_waits_for_torch(ctx):
    """The library issues its copies on its own non-blocking stream; torch (NCCL, cat, the caching allocator) works on
    torch's current stream.  Before the library touches a torch tensor, its stream must wait for torch's: without this
    the upload below could read the gathered array before torch.cat had written it (seen on 2 x B200: photon maps
    not byte-identical to the single-GPU trace)."""
    import torch
    dev = torch.cuda.current_device()
    torch.cuda.ExternalStream(ctx.stream, device=dev).wait_stream(torch.cuda.current_stream(dev))


def shoot_sharded(ctx, n_paths, max_depth=5, caustic=True, indirect=True, group=None):
    """Trace this rank's slice of [0, n_paths), then assemble both photon maps on every rank (all-gather).  Returns the
    stored counts (caustic, indirect).  The assembled arrays equal a single-GPU trace of the whole range byte for byte."""
    import torch
    import torch.distributed as dist
    from . import b2rt
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, count = path_range(n_paths, rank, world)
    ctx.photons_clear()
    ctx.photons_trace(first, count, max_depth=max_depth, caustic=caustic, indirect=indirect)
    totals = []
    for which in (b2rt.MAP_CAUSTIC, b2rt.MAP_INDIRECT):
        n, _ = ctx.photons_count(which)
        dev = torch.device("cuda", torch.cuda.current_device())
        local = torch.empty((n, 32), dtype=torch.uint8, device=dev)
        if n:
            _ctx_stream_waits_for_torch(ctx)        # `local` may reuse a block torch is still reading
            # the call synchronises the library stream before it returns, so torch may use `local` right away
            ctx._ck(ctx.lib.b2rt_photons_download(ctx.h, which, local.data_ptr(), n, b2rt.DEVICE))
        allp, counts = allgather_varlen(local, group)
        allp = allp.contiguous()
        _ctx_stream_waits_for_torch(ctx)            # all_gather + cat must have finished before the library copies
        on = (caustic if which == b2rt.MAP_CAUSTIC else indirect)
        ctx._ck(ctx.lib.b2rt_photons_upload(ctx.h, which, allp.data_ptr(), allp.shape[0], n_paths if on else 0, b2rt.DEVICE))
        totals.append(int(allp.shape[0]))
    return tuple(totals)


def render_sharded(ctx, cam, group=None, **kw):
    """Each rank renders its band; rank 0 returns the composited (yres, xres, 3) image, the others None."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    x0, x1, y0, y1 = band_window(cam.yres, cam.xres, rank, world)
    band = ctx.render_frame(cam, window=(x0, x1, y0, y1), **kw) if y1 > y0 else np.zeros((0, cam.xres, 3), np.float32)
    t = torch.from_numpy(band.reshape(-1, cam.xres * 3))
    if torch.cuda.is_available() and dist.get_backend(group) == "nccl":
        t = t.cuda()
    allb, _ = allgather_varlen(t, group)
    if rank != 0:
        return None
    return allb.cpu().numpy().reshape(cam.yres, cam.xres, 3)
