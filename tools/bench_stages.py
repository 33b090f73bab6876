"""Roofline fractions of the HBM-bound stages: K2 radix sort, K1-K4 LBVH build, K10/K11 grid build.
Algorithmic bytes per unit are the figures of SURVEY 8(d) / DESIGN.md section 3; time is device time (CUDA events on
the library stream); peak = MEASURED_PEAKS.json hbm_gbs (fallback 6650 GB/s).  Prints one JSON object.
python tools/bench_stages.py [sort_log2n ...]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

peak = 6650.0
pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
if os.path.exists(pk):
    peak = json.load(open(pk)).get("hbm_gbs", peak)
out = {"peak_gbs": peak, "sort": [], "lbvh_build": [], "grid_build": [], "film": []}


def med(f, reps=7):
    ts = []
    for _ in range(reps):
        ts.append(f())
    return float(np.median(ts))


# ---- K2: sort of (key32, value32) pairs, 4 passes: n * (4 + 16 * 4) = 68 B / element -------------------------------
with b2rt.Context(0, seed=1) as ctx:
    stream = torch.cuda.ExternalStream(ctx.stream, device=0)
    for lg in [int(a) for a in sys.argv[1:]] or [20, 24, 26]:
        n = 1 << lg
        with torch.cuda.stream(stream):
            src = torch.randint(0, 2 ** 31 - 1, (n,), dtype=torch.int32, device="cuda")
            keys, vals = torch.empty_like(src), torch.empty_like(src)
        ctx.sync()

        def one():
            with torch.cuda.stream(stream):
                keys.copy_(src)
                vals.copy_(torch.arange(n, dtype=torch.int32, device="cuda"))
            ctx.timer_start()
            ctx.sort_pairs_device(n, keys.data_ptr(), vals.data_ptr(), 32)
            return ctx.timer_stop()
        one()
        ms = med(one)
        ok = bool((keys[1:] >= keys[:-1]).all().item())
        gbs = n * 68 / (ms * 1e-3) / 1e9
        out["sort"].append({"n": n, "ms": ms, "Mkeys_per_s": n / ms / 1e3, "algorithmic_GBs": gbs, "frac_of_peak": gbs / peak,
                            "sorted": ok})
        del src, keys, vals

# ---- K1-K4: LBVH build, 436 B / primitive ------------------------------------------------------------
for ntris in (1_000_000, 8_000_000):
    P, idx = S.cfg2_mesh(ntris)
    with b2rt.Context(0, seed=1) as ctx:
        ctx.add_mesh(P, idx)
        infos = [ctx.build_accel() for _ in range(4)]
        ms = float(np.median([i.build_ms for i in infos[1:]]))
        gbs = len(idx) * 436 / (ms * 1e-3) / 1e9
        out["lbvh_build"].append({"n_prims": int(len(idx)), "build_ms": ms, "upload_ms": float(infos[-1].upload_ms),
                                  "first_build_ms": float(infos[0].build_ms), "Mprims_per_s": len(idx) / ms / 1e3,
                                  "algorithmic_GBs": gbs, "frac_of_peak": gbs / peak, "wide_nodes": int(infos[-1].n_wide_nodes)})
    del P, idx

# ---- K10/K11: hash grid over the photons of the Cfg3 scene, 188 B / photon + 16 B position copy ----------------------
with b2rt.Context(0, seed=3) as ctx:
    S.fill_scene(ctx, S.cornell_scene(spheres=True, wide=1.78))
    ctx.build_accel()
    for n_paths in (2_400_000, 24_000_000, 96_000_000):
        ctx.photons_clear()
        for first in range(0, n_paths, 24_000_000):
            ctx.photons_trace(first, min(24_000_000, n_paths - first))
        n = ctx.photons_count(b2rt.MAP_INDIRECT)[0]
        gs = [ctx.grid_build(b2rt.MAP_INDIRECT) for _ in range(4)]
        ms = float(np.median([g.build_ms for g in gs[1:]]))
        gbs = n * 204 / (ms * 1e-3) / 1e9
        out["grid_build"].append({"n_photons": int(n), "build_ms": ms, "Mphotons_per_s": n / ms / 1e3, "cells": list(gs[-1].dims),
                                  "algorithmic_GBs": gbs, "frac_of_peak": gbs / peak})
# ---- K13: film reconstruction, 20 B / sample (position + radiance) + 12 B / pixel; every sample is re-read by the
# (2h+1)^2 pixels whose support it may touch, out of L2 --------------------------------------------------------------
with b2rt.Context(0, seed=1) as ctx:
    stream = torch.cuda.ExternalStream(ctx.stream, device=0)
    for name, xres, yres, spp in (("box", 1920, 1080, 4), ("mitchell", 1920, 1080, 4), ("mitchell", 3840, 2160, 16),
                                  ("gaussian", 3840, 2160, 16), ("sinc", 1920, 1080, 4)):
        flt = b2rt.pbrt_filter(name)
        h = int(np.floor(flt.xwidth + 0.5))
        cells = (-h, -h, xres + h, yres + h)
        cw, ch = cells[2] - cells[0], cells[3] - cells[1]
        n = cw * ch * spp
        with torch.cuda.stream(stream):
            cx = torch.arange(cells[0], cells[2], device="cuda", dtype=torch.float32).view(1, cw, 1).expand(ch, cw, spp)
            cy = torch.arange(cells[1], cells[3], device="cuda", dtype=torch.float32).view(ch, 1, 1).expand(ch, cw, spp)
            xy = (torch.stack([cx, cy], -1) + torch.rand(ch, cw, spp, 2, device="cuda")).reshape(n, 2).contiguous()
            rgb = torch.rand(n, 3, device="cuda")
            img = torch.empty(yres, xres, 3, device="cuda")
        ctx.sync()

        def one():
            ctx.timer_start()
            ctx.film_reconstruct_device(flt, cells, spp, xy.data_ptr(), rgb.data_ptr(), (0, 0, xres, yres), img.data_ptr())
            return ctx.timer_stop()
        one()
        ms = med(one)
        gbs = (n * 20 + xres * yres * 12) / (ms * 1e-3) / 1e9
        out["film"].append({"filter": name, "xres": xres, "yres": yres, "spp": spp, "samples": n, "ms": ms,
                            "Msamples_per_s": n / ms / 1e3, "algorithmic_GBs": gbs, "frac_of_peak": gbs / peak,
                            "mean": float(img.mean().item())})
        del xy, rgb, img
print(json.dumps(out))
