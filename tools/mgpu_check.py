"""Multi-GPU check, launched with torchrun (one process per GPU, NCCL):
  * photon maps assembled by shoot_sharded (all-gather) are byte-identical to a single-GPU trace of the same range,
  * the banded frame composited on rank 0 is bit-identical to the single-GPU frame,
  * timings of the sharded photon frame (max over ranks).
torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py [n_paths]"""
import hashlib
import os
import sys
import time
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S, multigpu as M   # noqa: E402

n_paths = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
xres, yres = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (480, 270)
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = b2rt.Context(local, seed=3)
S.fill_scene(ctx, S.cornell_scene(wide=1.78))
info = ctx.build_accel()
cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, xres, yres)
kw = dict(spp=1, nused=64, max_dist=0.1)

for rep in range(3):
    dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
    nc, ni = M.shoot_sharded(ctx, n_paths)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    ctx.grid_build(0); ctx.grid_build(1)
    ctx.sync(); t2 = time.perf_counter()
    img = M.render_sharded(ctx, cam, **kw)
    torch.cuda.synchronize(); t3 = time.perf_counter()
    tt = torch.tensor([t1 - t0, t2 - t1, t3 - t2, t3 - t0], device="cuda", dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("rep %d world %d: shoot+allgather %.2f ms, grids %.2f ms, frame+composite %.2f ms, total %.2f ms (%d+%d photons)"
              % (rep, world, *(tt * 1e3).tolist(), nc, ni), flush=True)

# ---- identity with the single-GPU result (computed on every rank; only hashes are compared) -----------------------
hs = hashlib.sha256()
for which in (0, 1):
    hs.update(ctx.photons_download(which).tobytes())
sharded_hash = hs.hexdigest()
ctx.photons_clear()
ctx.photons_trace(0, n_paths)
hs = hashlib.sha256()
for which in (0, 1):
    hs.update(ctx.photons_download(which).tobytes())
single_hash = hs.hexdigest()
ctx.grid_build(0); ctx.grid_build(1)
full = ctx.render_frame(cam, **kw)
ok_img = True
if rank == 0:
    ok_img = bool(np.array_equal(full.view(np.uint32), img.view(np.uint32)))
# the same identity with a pixel filter wider than the pixel (every band re-traces its halo rows) and pbrt's
# low-discrepancy pixel pattern
ctx.set_filter(b2rt.pbrt_filter("mitchell")); ctx.set_sampler("lowdiscrepancy")
dist.barrier(); torch.cuda.synchronize(); t0 = time.perf_counter()
img_f = M.render_sharded(ctx, cam, **kw)
torch.cuda.synchronize(); t1 = time.perf_counter()
full_f = ctx.render_frame(cam, **kw)
ok_filtered = True
if rank == 0:
    ok_filtered = bool(np.array_equal(full_f.view(np.uint32), np.asarray(img_f).view(np.uint32)))
    print("mitchell + lowdiscrepancy banded frame: %.2f ms on rank 0" % ((t1 - t0) * 1e3), flush=True)
ctx.set_filter(None); ctx.set_sampler("random")
# the same identities with final gathering: sharded second photon pass (direct + radiance photons), replicated Lo, banded
# frame with the gather on (a smaller map keeps the single-GPU comparison frame short)
n_fg = min(n_paths, 400_000)
M.shoot_sharded(ctx, n_fg)
ctx.grid_build(0); ctx.grid_build(1)
M.shoot_sharded_fg(ctx, n_fg, nused=50, max_dist=0.1)
ctx.set_final_gather(True, 32, 10.0)
hs = hashlib.sha256()
for which in (2, 3):
    hs.update(ctx.photons_download(which).tobytes())
fg_sharded_hash = hs.hexdigest()
img_fg = M.render_sharded(ctx, cam, **kw)
ctx.set_final_gather(False)
ctx.photons_clear()
ctx.photons_trace(0, n_fg); ctx.photons_trace_fg(0, n_fg)
for which in (0, 1, 2):
    ctx.grid_build(which)
ctx.radiance_compute(50, 0.1)
hs = hashlib.sha256()
for which in (2, 3):
    hs.update(ctx.photons_download(which).tobytes())
fg_single_hash = hs.hexdigest()
ctx.set_final_gather(True, 32, 10.0)
full_fg = ctx.render_frame(cam, **kw)
ctx.set_final_gather(False)
ok_fg = fg_sharded_hash == fg_single_hash
if rank == 0:
    ok_fg = ok_fg and bool(np.array_equal(full_fg.view(np.uint32), np.asarray(img_fg).view(np.uint32)))
acc = hashlib.sha256(np.ascontiguousarray(ctx.accel_download()["wide_nodes"]).tobytes()).hexdigest()
objs = [None] * world
dist.all_gather_object(objs, (sharded_hash, single_hash, acc, ok_fg))
if rank == 0:
    same_photons = all(o[0] == o[1] == objs[0][0] for o in objs)
    same_bvh = all(o[2] == objs[0][2] for o in objs)
    print("photon maps identical to single-GPU on all ranks:", same_photons)
    print("BVH bytes identical across ranks:", same_bvh)
    print("banded frame identical to single-GPU frame:", ok_img)
    print("banded mitchell-filtered, low-discrepancy frame identical to single-GPU frame:", ok_filtered)
    same_fg = all(o[3] for o in objs)
    print("final gathering: direct / radiance photons (with Lo) and banded frame identical to single-GPU:", same_fg)
    assert same_photons and same_bvh and ok_img and ok_filtered and same_fg
ctx.close()
dist.destroy_process_group()
