"""Stage timings of the photon-mapping frame (BASELINE Cfg3 flavour): shoot, grid build, camera pass with kNN gather.
python tools/bench_photon.py [n_indirect] [n_caustic] [xres] [yres] [k] [reps]"""
import os
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

ni = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
nc = int(sys.argv[2]) if len(sys.argv) > 2 else 2_000_000
xres = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
yres = int(sys.argv[4]) if len(sys.argv) > 4 else 1080
k = int(sys.argv[5]) if len(sys.argv) > 5 else 64
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 3
desc = S.cornell_scene(spheres=True, boxes=False, wide=1.78)
with b2rt.Context(0, seed=3) as ctx:
    S.fill_scene(ctx, desc)
    ctx.build_accel()
    if os.environ.get("FILTER"):             # pbrt PixelFilter name: box / triangle / gaussian / mitchell / sinc
        ctx.set_filter(b2rt.pbrt_filter(os.environ["FILTER"]))
    cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, xres, yres)
    frame_out = ctx.pinned_empty((yres, xres, 3))
    for rep in range(reps):
        ctx.sync(); t0 = time.perf_counter()
        (pc, pi), (sc, si) = ctx.photons_shoot(nc, ni, batch_paths=4_000_000)
        ctx.sync(); t1 = time.perf_counter()
        g0 = ctx.grid_build(0); g1 = ctx.grid_build(1)
        ctx.sync(); t2 = time.perf_counter()
        img = ctx.render_frame(cam, spp=1, nused=k, max_dist=0.1, out=frame_out)
        ctx.sync(); t3 = time.perf_counter()
        imgd = ctx.render_frame(cam, spp=1, caustic=False, indirect=False)
        ctx.sync(); t4 = time.perf_counter()
        print("rep %d: shoot %.1f ms (%d+%d photons, %d/%d paths, %.1f Mpaths/s) | grids %.2f ms (build_ms %.2f+%.2f, cells %s %s c=%.4f) | "
              "frame %.1f ms | direct-only frame %.1f ms | mean %.4f"
              % (rep, (t1 - t0) * 1e3, sc, si, pc, pi, max(pc, pi) / (t1 - t0) / 1e6, (t2 - t1) * 1e3, g0.build_ms, g1.build_ms,
                 list(g0.dims), list(g1.dims), g1.cell_size, (t3 - t2) * 1e3, (t4 - t3) * 1e3, img.mean()))
