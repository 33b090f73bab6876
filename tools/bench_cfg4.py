"""BASELINE configs[3]: 8 M-triangle heightfield, 3840x2160, 16 spp, one camera ray + one shadow ray
per sample (matte surface, point light, direct lighting only) through b2rt_render_frame: camera rays, traversal,
shading and the film resolve all run on the GPU; the image is split into horizontal bands over the ranks and gathered
on rank 0 (multigpu.render_sharded).  Prints one JSON object on rank 0.
  python tools/bench_cfg4.py [tris] [xres] [yres] [spp]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_cfg4.py ..."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S, multigpu as M   # noqa: E402

tris = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
xres = int(sys.argv[2]) if len(sys.argv) > 2 else 3840
yres = int(sys.argv[3]) if len(sys.argv) > 3 else 2160
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 16
rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

P, idx = S.cfg2_mesh(tris)
ctx = b2rt.Context(local, seed=7)
mat = ctx.add_material(S.MATTE, (0.7, 0.7, 0.7), (0, 0, 0), 1.5)
ctx.add_mesh(P, idx, material=mat)
ctx.add_light(0, (30.0, 30.0, 30.0), p=(0.5, 3.0, 0.5))
info = ctx.build_accel()
if os.environ.get("FILTER"):                 # pbrt PixelFilter name: box / triangle / gaussian / mitchell / sinc
    ctx.set_filter(b2rt.pbrt_filter(os.environ["FILTER"]))
cam = b2rt.pbrt_camera((0, 1.2, 2.5), (0, 0, 0), (0, 1, 0), 45, xres, yres)
kw = dict(spp=spp, caustic=False, indirect=False, direct=True)
ctx.set_sampler("lowdiscrepancy")               # (0,2)-sequence jitter
skw = dict(kw, stripes=int(os.environ["STRIPES"])) if os.environ.get("STRIPES") else kw      # bands per rank (default: up to 8, at least 64 rows each)

frame_out = ctx.pinned_empty((yres, xres, 3))
times = []
img = None
for rep in range(4):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(); ctx.sync(); t0 = time.perf_counter()
    img = M.render_sharded(ctx, cam, **skw) if world > 1 else ctx.render_frame(cam, out=frame_out, **kw)
    torch.cuda.synchronize(); ctx.sync(); t1 = time.perf_counter()
    tt = torch.tensor([t1 - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    times.append(float(tt.item()) * 1e3)
ms = min(times[1:])
if rank == 0:
    samples = xres * yres * spp
    lit = float((img.sum(-1) > 0).mean())
    print(json.dumps({"workload": "Cfg4: %d triangles, %dx%d, %d spp (0,2)-sequence, camera ray + shadow ray per sample, interleaved bands over %d GPU(s)"
                                  % (len(idx), xres, yres, spp, world),
                      "n_gpus": world, "ms_per_frame": ms, "all_ms": times, "Msamples_per_s": samples / ms / 1e3,
                      "Mrays_per_s_upper": 2 * samples / ms / 1e3, "lit_pixel_fraction": lit,
                      "build_ms": info.build_ms, "upload_ms": info.upload_ms, "wide_nodes": info.n_wide_nodes,
                      "timing": "wall clock around the whole call incl. the device->host copy of the image, best of 3 after 1 warm-up, max over ranks"}),
          flush=True)
ctx.close()
if world > 1:
    dist.destroy_process_group()
