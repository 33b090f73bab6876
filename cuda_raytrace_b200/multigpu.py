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


def stripe_plan(yres, world, stripes=None):
    """Interleaved image stripes: the image rows are cut into stripes * world bands and band b belongs to rank
    b % world, so every rank samples the whole height of the image.  Returns the nb + 1 row boundaries.  Default stripe
    count: up to 8 per rank with bands of at least 64 rows (every band is one render call)."""
    if stripes is None:
        stripes = max(1, min(8, yres // (64 * world)))
    nb = max(1, int(stripes)) * world
    return [yres * b // nb for b in range(nb + 1)]


def stripe_rows(edges, world, rank):
    """[(y0, y1)] of the non-empty bands of `rank`, top to bottom - the order they are stacked in its buffer"""
    return [(edges[b], edges[b + 1]) for b in range(len(edges) - 1) if b % world == rank and edges[b + 1] > edges[b]]


def stripe_compose(parts, edges, world):
    """parts[r] = rank r's stacked bands (anything sliceable by rows); returns the slices in image order"""
    offs = [0] * world
    pieces = []
    for b in range(len(edges) - 1):
        r, n = b % world, edges[b + 1] - edges[b]
        if n:
            pieces.append(parts[r][offs[r]:offs[r] + n])
            offs[r] += n
    return pieces


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


def assemble_map(ctx, which, n_paths, group=None):
    """Replace this rank's slice of photon map `which` by the concatenation of all ranks' slices in rank order.
    Counts first (one all-gather, one host read), then the payload as ONE ncclAllGather of slices padded to the largest
    count, a device-side compaction of the padding (skipped when the counts are equal) and one copy that hands the array
    to the library.  Earlier versions: a list all-gather + cat through `allgather_varlen` (slow: per-rank copies in and
    out of a flat buffer), then one in-place broadcast per rank (8 collectives per map: 341 GB/s of bus bandwidth for 4 GB
    on 8 GPUs, 115 GB/s for 400 MB).  `B2RT_ASSEMBLE=broadcast` selects the broadcast version for A/B runs.
    Returns the total photon count."""
    import os
    import torch
    import torch.distributed as dist
    from . import b2rt
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    n, _ = ctx.photons_count(which)
    dev = torch.device("cuda", torch.cuda.current_device())
    mine = torch.tensor([n], dtype=torch.int64, device=dev)
    counts_t = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts_t, mine, group=group)
    counts = [int(c) for c in counts_t.tolist()]
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    total = int(offs[-1])
    if os.environ.get("B2RT_ASSEMBLE") == "broadcast":
        out = torch.empty((max(total, 1), 32), dtype=torch.uint8, device=dev)
        if n:
            _ctx_stream_waits_for_torch(ctx)        # `out` may reuse a block torch is still reading
            # synchronous on return, so torch may broadcast from the slice right away
            ctx._ck(ctx.lib.b2rt_photons_download(ctx.h, which, out[int(offs[rank]):].data_ptr(), n, b2rt.DEVICE))
        works = [dist.broadcast(out[int(offs[r]):int(offs[r + 1])], src=dist.get_global_rank(group, r) if group else r,
                                group=group, async_op=True) for r in range(world) if counts[r]]
        for w in works:
            w.wait()
    else:
        mx = max(max(counts), 1)
        stage = torch.empty((world, mx, 32), dtype=torch.uint8, device=dev)
        inp = torch.empty((mx, 32), dtype=torch.uint8, device=dev)
        if n:
            _ctx_stream_waits_for_torch(ctx)        # `inp` may reuse a block torch is still reading
            ctx._ck(ctx.lib.b2rt_photons_download(ctx.h, which, inp.data_ptr(), n, b2rt.DEVICE))
        dist.all_gather_into_tensor(stage.view(-1), inp.view(-1), group=group)
        if all(c == mx for c in counts):
            out = stage.view(world * mx, 32)
        else:
            out = torch.cat([stage[r, :counts[r]] for r in range(world)]) if total else stage.view(-1, 32)[:1]
    _ctx_stream_waits_for_torch(ctx)                # the collective must have finished before the library copies
    ctx._ck(ctx.lib.b2rt_photons_upload(ctx.h, which, out.data_ptr(), total, n_paths, b2rt.DEVICE))
    return total


def shoot_sharded(ctx, n_paths, max_depth=5, caustic=True, indirect=True, group=None):
    """Trace this rank's slice of [0, n_paths), then assemble both photon maps on every rank.  Returns the stored counts
    (caustic, indirect).  The assembled arrays equal a single-GPU trace of the whole range byte for byte."""
    import torch.distributed as dist
    from . import b2rt
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, count = path_range(n_paths, rank, world)
    ctx.photons_clear()
    ctx.photons_trace(first, count, max_depth=max_depth, caustic=caustic, indirect=indirect)
    return (assemble_map(ctx, b2rt.MAP_CAUSTIC, n_paths if caustic else 0, group),
            assemble_map(ctx, b2rt.MAP_INDIRECT, n_paths if indirect else 0, group))


_pinned_frames = {}


def shoot_sharded_fg(ctx, n_paths, max_depth=5, nused=50, max_dist=0.1, group=None):
    """Final gathering's second photon pass, sharded like the first: this rank re-traces its slice of [0, n_paths) for the
    direct and radiance photons (b2rt_photons_trace_fg), both maps are assembled on every rank (rank-order concatenation
    == the single-GPU arrays byte for byte), then every rank builds the grids and computes the radiance photons' Lo itself
    (replicated, deterministic: the same bytes everywhere, and no exchange of per-photon results).  Call after
    shoot_sharded + grid_build of maps 0 and 1.  Returns the stored counts (direct, radiance)."""
    import torch.distributed as dist
    from . import b2rt
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    first, count = path_range(n_paths, rank, world)
    ctx.photons_trace_fg(first, count, max_depth=max_depth)
    nd = assemble_map(ctx, b2rt.MAP_DIRECT, n_paths, group)
    nr = assemble_map(ctx, b2rt.MAP_RADIANCE, n_paths, group)
    if nd:
        ctx.grid_build(b2rt.MAP_DIRECT)
    ctx.radiance_compute(nused, max_dist)
    return nd, nr


def render_sharded(ctx, cam, group=None, **kw):
    """Each rank renders its bands (interleaved stripes, `stripes=` per rank, default up to 8 with at least 64 rows per band); rank 0 returns the composited
    (yres, xres, 3) image, the others None.  With NCCL the
    returned array is a view of a pinned buffer that the next call with the same resolution overwrites.
    NCCL: the band is rendered straight into a CUDA tensor, the bands are gathered on rank 0's GPU and the finished image
    crosses PCIe once, into pinned memory (the first version went band -> pageable host -> device -> all-gather -> host
    and made the 4K frame slower on two GPUs than on one).  gloo (CPU tests): host arrays all the way."""
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    x0, x1, y0, y1 = band_window(cam.yres, cam.xres, rank, world)
    stripes = kw.pop("stripes", None)
    if not (torch.cuda.is_available() and dist.get_backend(group) == "nccl"):
        band = ctx.render_frame(cam, window=(x0, x1, y0, y1), **kw) if y1 > y0 else np.zeros((0, cam.xres, 3), np.float32)
        allb, _ = allgather_varlen(torch.from_numpy(band.reshape(-1, cam.xres * 3)), group)
        return allb.numpy().reshape(cam.yres, cam.xres, 3) if rank == 0 else None
    dev = torch.device("cuda", torch.cuda.current_device())
    # interleaved stripes: the image is cut into stripes * world bands and rank r renders bands r, r + world, ... so
    # every rank samples the whole height of the image (a rank whose single band was all terrain took 1.7x the mean
    # band time on the 8 M-triangle 4K frame; SURVEY 8(e) asks for interleaved tiles for this reason)
    edges = stripe_plan(cam.yres, world, stripes)
    pad = max(max(sum(y1 - y0 for y0, y1 in stripe_rows(edges, world, r)) for r in range(world)), 1)
    band = torch.empty((pad, cam.xres * 3), dtype=torch.float32, device=dev)
    _ctx_stream_waits_for_torch(ctx)                # `band` may reuse a block torch is still reading
    at = 0
    for by0, by1 in stripe_rows(edges, world, rank):
        ctx.render_frame_device(cam, band[at:].data_ptr(), window=(0, cam.xres, by0, by1), **kw)   # synchronous on return
        at += by1 - by0
    parts = [torch.empty_like(band) for _ in range(world)] if rank == 0 else None
    # `dst` is a GLOBAL rank: translate the group's rank 0 (a sub-group need not contain global rank 0)
    dist.gather(band, parts, dst=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    if rank != 0:
        return None
    img = torch.cat(stripe_compose(parts, edges, world), 0)
    key = (cam.yres, cam.xres)
    if key not in _pinned_frames:
        _pinned_frames[key] = torch.empty((cam.yres, cam.xres * 3), dtype=torch.float32).pin_memory()
    host = _pinned_frames[key]
    host.copy_(img, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return host.numpy().reshape(cam.yres, cam.xres, 3)
