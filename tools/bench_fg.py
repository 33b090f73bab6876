"""Stage timings of a frame with final gathering (pbrt photonmap, finalgather true): second photon pass (direct + radiance
photons), ComputeRadiance, camera pass with the gather.  Cornell + glass / mirror spheres.
python tools/bench_fg.py [n_indirect] [n_caustic] [xres] [yres] [gather_samples] [reps]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

ni = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
nc = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
xres = int(sys.argv[3]) if len(sys.argv) > 3 else 960
yres = int(sys.argv[4]) if len(sys.argv) > 4 else 540
gs = int(sys.argv[5]) if len(sys.argv) > 5 else 32
reps = int(sys.argv[6]) if len(sys.argv) > 6 else 2
desc = S.cornell_scene(spheres=True, boxes=False, wide=1.78)
with b2rt.Context(0, seed=3) as ctx:
    S.fill_scene(ctx, desc)
    ctx.build_accel()
    cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, xres, yres)
    for rep in range(reps):
        ctx.photons_clear()
        ctx.sync(); t0 = time.perf_counter()
        (pc, pi), (sc, si) = ctx.photons_shoot(nc, ni, batch_paths=4_000_000)
        ctx.sync(); t1 = time.perf_counter()
        nd, nr = ctx.photons_trace_fg(0, pi)
        ctx.sync(); t2 = time.perf_counter()
        for m in (0, 1, 2):
            ctx.grid_build(m)
        ctx.sync(); t3 = time.perf_counter()
        ctx.radiance_compute(50, 0.1)
        ctx.sync(); t4 = time.perf_counter()
        ctx.set_final_gather(True, gs, 10.0)
        n0 = ctx.launch_count(reset=True)
        img = ctx.render_frame(cam, spp=1, nused=50, max_dist=0.1)
        ctx.sync(); t5 = time.perf_counter()
        launches = ctx.launch_count(reset=True)
        ctx.set_final_gather(False)
        img0 = ctx.render_frame(cam, spp=1, nused=50, max_dist=0.1)
        ctx.sync(); t6 = time.perf_counter()
        rays = xres * yres * 2 * max(1, gs // 2)
        print(json.dumps({"rep": rep, "photons": {"caustic": sc, "indirect": si, "direct": nd, "radiance": nr, "paths": pi},
                          "shoot_ms": (t1 - t0) * 1e3, "fg_photon_pass_ms": (t2 - t1) * 1e3, "grids_ms": (t3 - t2) * 1e3,
                          "radiance_compute_ms": (t4 - t3) * 1e3, "fg_frame_ms": (t5 - t4) * 1e3, "fg_frame_launches": launches,
                          "gather_rays_upper_bound": rays, "gather_mrays_per_s_lower_bound": rays / (t5 - t4) / 1e6,
                          "plain_frame_ms": (t6 - t5) * 1e3, "mean_fg": float(img.mean()), "mean_plain": float(img0.mean()),
                          "res": [xres, yres], "gather_samples": gs}), flush=True)
