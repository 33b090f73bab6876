#!/usr/bin/env python
"""bench.py -- the driver contract (one JSON line on stdout, rank 0).

Workload (N=1): BASELINE.json configs[1] = "Synthetic 1M-triangle procedural mesh, 1920x1080 primary+shadow
rays 1 spp (intersection throughput)" -- SURVEY.md 8(d) Cfg2.  One step = one pass of the hot path over one
frame's rays: closest-hit over the 2 073 600 primary rays, then any-hit over the shadow rays of every primary
hit (the reference's SimpleRenderer frame, cuda_render/simple_render/simplerender.cu:22-28,47-78).

  value      Mrays/s = (primary + shadow rays of all ranks) / max-over-ranks device time, rays resident in HBM
  e2e        same metric through the C ABI with HOST (pinned) buffers: b2rt_trace_closest/any(..., B2RT_HOST)
             copy the rays in and the hits / occlusion flags out inside the timed region
  roofline   dominant kernel = trace_spec_kernel<closest> (persistent ray-pool traversal); algorithmic bytes = 48 B/ray (32 in + 16 out)
  cpu_baseline  the reference's own CPU path (pbrt-v2 BVHAccel::Intersect/IntersectP from oracle/_ref, built from
             /root/reference in the build container) on the same rays, all host threads; falls back to nothing
             else -- if oracle/_ref is absent the C port's exhaustive loop is timed on a small sample ("port")

N>1 (torchrun): weak scaling, no data-path collective -- rank r traces sample r of every pixel (its own
1920x1080 ray set with a rank-specific sub-pixel offset) against its own replicated, deterministic BVH.

--impl reference: times the reference's CPU implementation of the same step on the host cores (rank 0 only).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints "NCCL version ..." from C at
# communicator creation, seen at N = 2), so keep a private copy of the real stdout for the JSON line and send
# everything else that targets fd 1 to stderr.
_JSON_OUT = None


def grab_stdout():
    global _JSON_OUT
    if _JSON_OUT is None:
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    grab_stdout()
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()

_TV = os.environ.get("B2RT_TRACE_VARIANT", "2")
TRACE_KERNEL = {"0": "trace_kernel<closest>", "1": "trace_persistent_kernel<closest>"}.get(_TV, "trace_spec_kernel<closest>")
# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the default closest-hit kernel on this exact workload,
# from the ncu --set full capture of `bench.py --steps 2 --warmup 3` (profiles/r01h_ncu_trace.md, re-captured r01m and
# r07a, profiles/r07a_ncu_trace_full_summary.md; r02m had 135-136 + 26): 135 MB read + 24.1 MB written, against 99.5 MB of compulsory ray / hit bytes (the difference is the cold 52 MB scene).
TRAFFIC_BYTES = 159.1e6 if _TV == "2" else None
TRAFFIC_SOURCE = "ncu --set full, r07a capture (profiles/r07a_ncu_trace_full_summary.md): dram read 135 MB + write 24.1 MB per launch"
METRIC = "Mrays/s (closest-hit primary + any-hit shadow rays, 1M-triangle scene, 1920x1080, 1 spp)"
LIGHT = (0.5, 3.0, 0.5)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--tris", type=int, default=1_000_000)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--cpu-sample", type=int, default=0, help="rays in the cpu_baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-photon-extra", action="store_true", help="skip the Cfg3 photon-frame stage timings")
    return ap.parse_args()


def make_workload(args, rank):
    from cuda_raytrace_b200 import scenes as S
    P, idx = S.cfg2_mesh(args.tris)
    if rank == 0:
        rays = S.pinhole_rays(args.width, args.height)          # pixel centres
    else:
        rays = S.pinhole_rays(args.width, args.height, spp=2, seed=1000 + rank)[::2].copy()  # jittered sample
    return P, idx, rays


class ClockSampler:
    """nvidia-smi clocks line of /opt/skills/guides/B200_PROFILING.md, sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def hit_parity(gpu_hits, gpu_occ, ref):
    """The step's GPU results against what the reference's own BVHAccel returned for the same rays (the cpu_baseline leg
    computes them anyway).  Bars (north star): hit distance bit-identical, primitive id identical except where two
    primitives share the winning t exactly (a ray through a shared edge: pbrt keeps whichever its tree meets last, the
    GPU the lower id) -- those are counted as ties, not mismatches."""
    s1, s2 = ref["primary_stride"], ref["shadow_stride"]
    gt = np.ascontiguousarray(gpu_hits[::s1, 0]).view(np.uint32)
    gp = np.ascontiguousarray(gpu_hits[::s1, 1]).view(np.int32)
    rt, rp = ref["t"].view(np.uint32), ref["prim"]
    t_bad = gt != rt
    p_bad = gp != rp
    go = np.ascontiguousarray(gpu_occ[::s2])
    out = _parity_counts(rt, rp, gt, gp, t_bad, p_bad, go, ref)
    bad = np.nonzero(t_bad)[0]
    if len(bad) and ref.get("brute") is not None:
        # who is right where the two disagree: the exhaustive loop over ALL primitives with pbrt's own Shape::Intersect
        # (renderers/aggregatetest.cpp:93-102) for exactly those rays.  pbrt's BVHAccel prunes with an unpadded box test
        # (accelerators/bvh.cpp:126-148) and can drop a hit that lies on a box face; the GPU traversal pads its boxes.
        bt, bp = ref["brute"](bad[:4096] * s1)
        bt = bt.view(np.uint32)
        out["disagreements_resolved_by_exhaustive_loop"] = {
            "rays": int(len(bt)), "gpu_equals_exhaustive": int((gt[bad[:4096]] == bt).sum()),
            "bvhaccel_equals_exhaustive": int((rt[bad[:4096]] == bt).sum()),
            "first": [{"ray": int(i * s1), "gpu_t": float(gt[i:i + 1].view(np.float32)[0]), "gpu_prim": int(gp[i]),
                       "bvhaccel_t": float(rt[i:i + 1].view(np.float32)[0]), "bvhaccel_prim": int(rp[i]),
                       "exhaustive_t": float(bt[j:j + 1].view(np.float32)[0]), "exhaustive_prim": int(bp[j])}
                      for j, i in enumerate(bad[:8])]}
        out["ok"] = bool(out["occluded_mismatch"] == 0 and (gt[bad[:4096]] == bt).all() and len(bad) <= 4096)
    return out


def _parity_counts(rt, rp, gt, gp, t_bad, p_bad, go, ref):
    return {"oracle": "pbrt-v2 BVHAccel::Intersect / IntersectP (oracle/_ref, the reference's own CPU path)",
            "rays": int(len(rt) + len(go)), "primary_rays": int(len(rt)), "shadow_rays": int(len(go)),
            "hits": int((rp >= 0).sum()), "t_mismatch": int(t_bad.sum()), "prim_mismatch": int((p_bad & t_bad).sum()),
            "ties": int((p_bad & ~t_bad).sum()), "occluded_mismatch": int((go != ref["occ"]).sum()),
            "ok": bool(not t_bad.any() and not (go != ref["occ"]).any())}


def cpu_reference_step(P, idx, rays, shadow, nthreads, sample=None):
    """The reference's CPU path on (a sample of) one step.  Returns dict(value Mrays/s, kind, cores, sample) and, for
    the reference kind, the hits it computed (for hit_parity)."""
    from oracle import oracle as O
    if O.ref_available():
        rs = O.RefScene()
        rs.add_mesh(P, idx)
        build_s = rs.build()
        nth = nthreads or rs.num_cores()
        st1 = 1 if sample is None else max(1, len(rays) // sample)
        st2 = 1 if sample is None else max(1, len(shadow) // sample)
        r1, r2 = rays[::st1], shadow[::st2]
        rs.intersect(r1[: 20000], nth)                                   # warm the page cache / threads
        t, prim, _, s1 = rs.intersect(r1, nth)
        occ, s2 = rs.occluded(r2, nth)
        n = len(r1) + len(r2)

        def brute(ray_idx):          # keeps the pbrt scene alive for hit_parity; closed there
            bt, bp, _ = rs.intersect_brute(rays[ray_idx], nth)
            return bt, bp
        return {"value": n / (s1 + s2) / 1e6, "unit": "Mrays/s", "cores": nth, "kind": "reference",
                "sample": "%d primary (BVHAccel::Intersect) + %d shadow (IntersectP) rays of the step, pbrt-v2 "
                          "BVHAccel built in %.2f s (excluded); ranges of 4096 rays handed out by pbrt's own task "
                          "queue (EnqueueTasks)" % (len(r1), len(r2), build_s),
                "closest_mrays": len(r1) / s1 / 1e6, "any_mrays": len(r2) / s2 / 1e6, "seconds": s1 + s2,
                "_hits": {"t": t, "prim": prim, "occ": occ, "primary_stride": st1, "shadow_stride": st2, "brute": brute,
                          "close": rs.close}}
    # port: exhaustive loop over all prims -- O(n_prims) per ray, so only a tiny sample
    ps = O.PortScene()
    ps.add_mesh(P, idx)
    k = 64 if sample is None else max(16, min(sample, 256))
    r1 = rays[:: max(1, len(rays) // k)]
    nth = nthreads or os.cpu_count()
    t0 = time.perf_counter()
    ps.intersect(r1, nth)
    s = time.perf_counter() - t0
    return {"value": len(r1) / s / 1e6, "unit": "Mrays/s", "cores": nth, "kind": "port",
            "sample": "%d primary rays, exhaustive loop over all prims (no CPU BVH in the port)" % len(r1),
            "seconds": s}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU implementation of the step, host cores, rank 0 only."""
    if rank != 0:
        return
    P, idx, rays = make_workload(args, 0)
    from oracle import oracle as O
    if O.ref_available():
        rs = O.RefScene()
        rs.add_mesh(P, idx)
        build_s = rs.build()
        nth = rs.num_cores()
        t, prim, _, _ = rs.intersect(rays, nth)
        from cuda_raytrace_b200 import scenes as S
        shadow = S.shadow_rays(rays, t, prim, LIGHT)
        for _ in range(args.warmup):
            rs.intersect(rays, nth); rs.occluded(shadow, nth)
        secs = []
        for _ in range(args.steps):
            _, _, _, s1 = rs.intersect(rays, nth)
            _, s2 = rs.occluded(shadow, nth)
            secs.append(s1 + s2)
        rs.close()
        total = sum(secs)
        nrays = len(rays) + len(shadow)
        v = nrays * args.steps / total / 1e6
        kind, sample = "reference", ("full step: %d primary + %d shadow rays through pbrt-v2 BVHAccel "
                                     "(built in %.2f s, excluded)" % (len(rays), len(shadow), build_s))
    else:
        cb = cpu_reference_step(P, idx, rays, rays, None, sample=64)
        v, nth, kind, sample, total = cb["value"], cb["cores"], cb["kind"], cb["sample"], cb["seconds"]
        args.steps = 1
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": "Mrays/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, len(idx)),
            "cpu_baseline": {"value": v, "unit": "Mrays/s", "cores": nth, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def workload_config(args, ntris):
    return {"workload": "Cfg2: %d-triangle fbm heightfield (seed 1234, shuffled), %dx%d pinhole primary rays 1 spp + "
                        "one shadow ray per hit to a point light" % (ntris, args.width, args.height),
            "cache": "L2 flushed between timed steps (256 MiB write); step working set ~220 MB > 126 MB L2",
            "sharding": "rank r traces sample r of every pixel against its own replicated BVH (no collective)"}


def photon_frame_extra(n_gpus, peak_gbs, cpu_baseline=True):
    """Not part of `value`: the other two clauses of the metric, "photons traced+gathered/s" and "ms/frame", on BASELINE
    configs[2] (Cornell + glass / mirror spheres, ~10 M stored photons, kNN k = 64, 1920x1080, 1 spp).  ONE process drives
    all n_gpus through the multi-GPU layer of the C ABI (b2rt_multi_*, csrc/multi.cu): photon paths sharded by index range
    with the exchange fused into the compaction kernel (peer stores over NVLink), replicated grids, 64 x 64 image tiles
    composited on GPU 0.  Stage times are DEVICE times: CUDA events on every GPU's library stream around the stage, max
    over the GPUs; median of 5 after 1 warm-up.  SHA-256 of both photon maps and of the frame are emitted so that the
    records of different GPU counts can be compared (the bytes must not depend on N)."""
    import hashlib
    from cuda_raytrace_b200 import b2rt, scenes as S
    n_paths = 24_000_000      # yields ~12 M indirect + ~0.53 M caustic photons in this scene
    k, maxdist = 64, 0.1
    desc = S.cornell_scene(spheres=True, wide=1.78)
    cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 1920, 1080)
    out = {}
    with b2rt.MultiContext(n_gpus, seed=3) as m:
        m.fill(lambda c: S.fill_scene(c, desc))
        m.build_accel()
        frame = m.ctxs[0].pinned_empty((1080, 1920, 3))

        def timed(fn):
            for c in m.ctxs:
                c.timer_start()
            r = fn()
            return r, max(c.timer_stop() for c in m.ctxs)
        rows = []
        for rep in range(6):
            m.photons_clear()
            (nc, ni), t_shoot = timed(lambda: m.photons_trace(0, n_paths))
            _, t_grid = timed(lambda: (m.grid_build(0), m.grid_build(1)))
            _, t_cam = timed(lambda: m.render_frame(cam, spp=1, nused=k, max_dist=maxdist, out=frame))
            if rep:
                rows.append((t_shoot, t_grid, t_cam, t_shoot + t_grid + t_cam))
        med = [float(np.median([r[j] for r in rows])) for j in range(4)]
        st = m.stats(reset=True)
        sha = {"caustic_map": hashlib.sha256(m.ctxs[-1].photons_download(0).tobytes()).hexdigest(),
               "indirect_map": hashlib.sha256(m.ctxs[-1].photons_download(1).tobytes()).hexdigest(),
               "frame": hashlib.sha256(np.ascontiguousarray(frame).tobytes()).hexdigest()}
        g0, g1 = m.ctxs[0].grid_info(0), m.ctxs[0].grid_info(1)
        nq = 1920 * 1080          # one diffuse leaf per pixel is the floor; specular pixels add leaves
        grid_bytes = 204.0 * (nc + ni)                                  # 188 B / photon + the 16-B position copy
        gather_bytes = 16.0 * (nc + ni) + 2 * nq * (32.0 * k + 60.0)     # every position once + k records and 60 B per query and map
        out = {"workload": "Cfg3: Cornell + glass/mirror spheres, %d emitted paths -> %d caustic + %d indirect photons, kNN k=%d, "
                           "maxdist %g, 1920x1080 1 spp" % (n_paths, nc, ni, k, maxdist),
               "api": "b2rt_multi_photons_trace / b2rt_multi_photon_grid_build / b2rt_multi_render_frame (C ABI, one process, %d GPU%s)"
                      % (n_gpus, "" if n_gpus == 1 else "s"),
               "multi": m.info(),
               "shoot_exchange_ms": med[0], "grid_build_ms": med[1], "camera_pass_ms": med[2], "frame_ms": med[3],
               "photons_traced_and_gathered_per_s": (nc + ni) / (med[3] * 1e-3), "paths_per_s": n_paths / (med[0] * 1e-3),
               "gather_queries_per_s": 2 * nq / (med[2] * 1e-3),
               "exchange": {"bytes_held_per_gpu": st["exchange_bytes"] // max(st["passes"], 1),
                            "bus_GBs": ((n_gpus - 1) / n_gpus * st["exchange_bytes"] / max(st["passes"], 1)) /
                                       max(st["exchange_ms"] / max(st["passes"], 1) * 1e-3, 1e-9) / 1e9 if n_gpus > 1 else None,
                            "host_wall_ms_per_pass": {kk: st[kk] / max(st["passes"], 1) for kk in ("trace_ms", "reserve_ms", "exchange_ms")}},
               "roofline": {"grid_build": {"bound": "hbm", "algorithmic_bytes": grid_bytes, "achieved": grid_bytes / (med[1] * 1e-3) / 1e9,
                                           "peak": peak_gbs, "unit": "GB/s", "frac": grid_bytes / (med[1] * 1e-3) / 1e9 / peak_gbs,
                                           "note": "replicated on every GPU: does not shrink with N"},
                            "camera_pass_gather": {"bound": "issue slots / L2 (not HBM)", "algorithmic_bytes": gather_bytes,
                                                   "achieved": gather_bytes / (med[2] * 1e-3) / 1e9 * n_gpus ** 0, "peak": peak_gbs, "unit": "GB/s",
                                                   "frac": gather_bytes / (med[2] * 1e-3) / 1e9 / (peak_gbs * n_gpus),
                                                   "note": "16 N + (32 k + 60) Q per map; the time also holds the camera rays and direct lighting"}},
               "sha256": sha, "grid_cells": [list(g0.dims), list(g1.dims)],
               "timing": "CUDA events on each GPU's library stream around the stage, max over GPUs, median of 5 after 1 warm-up"}
        if cpu_baseline:
            # the reference's CPU path for the same clauses, bounded sample: PhotonIntegrator::Preprocess (photons/s, kd-tree
            # build included) and KdTree lookup + LPhoton (calls/s) on the GPU's own photons
            from oracle import oracle as O
            if O.ref_available():
                nth = os.cpu_count() or 1
                rp = O.RefPhotonScene()
                S.fill_scene(rp, desc)
                t0 = time.perf_counter()
                pc_, pi_, paths_ = rp.shoot(100000, 4000000, ncores=0)
                t_pre = time.perf_counter() - t0
                sub = m.ctxs[0].photons_download(1)[:4_000_000]
                nqs = 2_000_000
                rays = S.pinhole_rays(1920, 1080, eye=(0, 1, 3.4), look=(0, 1, 0), fov=40)
                sf = m.ctxs[0].surface_at(rays)
                sf = sf[sf["prim"] >= 0][:nqs]
                q = np.zeros(len(sf), b2rt.QUERY_DTYPE)
                q["p"], q["n"], q["kd"] = sf["p"], sf["n"], 0.5
                _, t_build, t_look = rp.lphoton_timed(O.photons_to9(sub), n_paths * len(sub) // max(ni, 1), q, k, maxdist * maxdist, nth)
                out["cpu_baseline"] = {
                    "kind": "reference", "cores": nth,
                    "preprocess": {"value": (len(pc_) + len(pi_)) / t_pre, "unit": "photons stored/s",
                                   "sample": "pbrt-v2 PhotonIntegrator::Preprocess, 100 k caustic + 4 M indirect photons wanted (%d + %d stored from "
                                             "%d paths), kd-tree builds included, %.2f s" % (len(pc_), len(pi_), max(paths_), t_pre)},
                    "lphoton": {"value": len(q) / t_look, "unit": "estimates/s",
                                "sample": "KdTree::Lookup + LPhoton, k = %d, %d visible points against %d of the GPU's indirect photons; "
                                          "kd-tree build %.2f s (excluded), lookups %.2f s" % (k, len(q), len(sub), t_build, t_look)}}
    return out


def main():
    grab_stdout()
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return run_reference(args, rank, world)

    import torch
    import torch.distributed as dist
    from cuda_raytrace_b200 import b2rt, scenes as S
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the b2rt path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    P, idx, rays = make_workload(args, rank)
    n1 = len(rays)
    ctx = b2rt.Context(local_rank, seed=1)
    ctx.add_mesh(P, idx)
    info = ctx.build_accel()
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)

    # primary hits once (host path) -> shadow rays, the simple renderer's second ray type
    hits0 = ctx.trace_closest(rays)
    shadow = S.shadow_rays(rays, hits0["t"], hits0["prim"], LIGHT)
    n2 = len(shadow)

    with torch.cuda.stream(stream):
        d_rays = torch.from_numpy(rays).cuda(non_blocking=False)
        d_shadow = torch.from_numpy(shadow).cuda(non_blocking=False)
        d_hits = torch.empty((n1, 4), dtype=torch.float32, device="cuda")
        d_occ = torch.empty(n2, dtype=torch.uint8, device="cuda")
        flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ctx.sync()

    def step_device():
        ctx.trace_closest_device(n1, d_rays.data_ptr(), d_hits.data_ptr())
        ctx.trace_any_device(n2, d_shadow.data_ptr(), d_occ.data_ptr())

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def flush_l2():
        with torch.cuda.stream(stream):
            flush.fill_(1)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    for _ in range(max(args.warmup, 3)):
        flush_l2(); step_device()
    ctx.sync()
    # a step is < 1 ms, nvidia-smi samples every 100 ms: keep the SAME load running (untimed) before and after
    # the timed steps so that the clock samples bracket the timed region under identical load
    def soak(seconds, min_samples):
        t_end = time.perf_counter() + seconds
        while time.perf_counter() < t_end or (rank == 0 and sampler.proc and len(sampler.lines) < min_samples
                                              and time.perf_counter() < t_end + 5.0):
            for _ in range(20):
                flush_l2(); step_device()
            ctx.sync()
    soak(1.0, 3)

    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    ctx.launch_count(reset=True)
    barrier()
    wall0 = time.perf_counter()
    for k in range(args.steps):
        flush_l2()
        ev[k][0].record(stream)
        ctx.trace_closest_device(n1, d_rays.data_ptr(), d_hits.data_ptr())
        ev[k][1].record(stream)
        ctx.trace_any_device(n2, d_shadow.data_ptr(), d_occ.data_ptr())
        ev[k][2].record(stream)
    ctx.sync()
    barrier()
    wall = time.perf_counter() - wall0
    launches = ctx.launch_count()
    soak(0.4, 0)
    clocks = sampler.stop() if rank == 0 else None
    ms_closest = sum(e[0].elapsed_time(e[1]) for e in ev)
    ms_any = sum(e[1].elapsed_time(e[2]) for e in ev)
    ms_total = ms_closest + ms_any

    # ---- extra (not part of `value`): incoherent closest-hit rays, the aggregatetest distribution ----------------
    inc = S.incoherent_rays(n1, info.bounds[0:3], info.bounds[3:6], seed=99 + rank)
    with torch.cuda.stream(stream):
        d_inc = torch.from_numpy(inc).cuda()
    e_inc = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    for _ in range(2):
        ctx.trace_closest_device(n1, d_inc.data_ptr(), d_hits.data_ptr())
    flush_l2()
    e_inc[0].record(stream)
    for _ in range(5):
        ctx.trace_closest_device(n1, d_inc.data_ptr(), d_hits.data_ptr())
    e_inc[1].record(stream)
    ctx.sync()
    inc_mrays = 5 * n1 / (e_inc[0].elapsed_time(e_inc[1]) * 1e-3) / 1e6
    del d_inc
    ctx.trace_closest_device(n1, d_rays.data_ptr(), d_hits.data_ptr())   # restore d_hits for the parity guard
    ctx.sync()

    # ---- e2e: the same step through the C ABI with host (pinned) buffers --------------------------------------
    h_rays = torch.from_numpy(rays).pin_memory()
    h_shadow = torch.from_numpy(shadow).pin_memory()
    h_hits = torch.empty((n1, 4), dtype=torch.float32).pin_memory()
    h_occ = torch.empty(n2, dtype=torch.uint8).pin_memory()
    lib, H = ctx.lib, ctx.h

    def step_host():
        rc = lib.b2rt_trace_closest(H, n1, h_rays.data_ptr(), h_hits.data_ptr(), b2rt.HOST)
        rc |= lib.b2rt_trace_any(H, n2, h_shadow.data_ptr(), h_occ.data_ptr(), b2rt.HOST)
        if rc:
            raise RuntimeError(lib.b2rt_last_error(H).decode())

    for _ in range(3):
        step_host()
    e2e_steps = args.steps
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        step_host()                      # each call ends with a stream sync (results are in host memory)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    barrier()
    # parity guard on the bench inputs themselves: device and host paths must agree bit for bit
    assert np.array_equal(h_hits.numpy().view(np.uint32), d_hits.cpu().numpy().view(np.uint32))
    gpu_hits_np, gpu_occ_np = h_hits.numpy().copy(), h_occ.numpy().copy()      # the timed step's own results (rank's rays)

    times = torch.tensor([ms_total, ms_closest, ms_any, e2e_s * 1e3], dtype=torch.float64, device="cuda")
    counts = torch.tensor([n1, n2], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    ms_total_m, ms_closest_m, ms_any_m, e2e_ms_m = times.tolist()
    N1, N2 = counts.tolist()

    ctx.close()
    # The photon-frame clauses run in ONE process over all N GPUs (the library's own multi-GPU layer), so the other ranks
    # are done: they leave (and release their GPUs) before rank 0 starts it.
    del d_rays, d_shadow, d_hits, d_occ, flush
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
        if rank != 0:
            return
    if rank == 0:
        peaks = {}
        pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk_path):
            peaks = json.load(open(pk_path))
        peak = peaks.get("hbm_gbs", 6650.0)
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        extra = None
        if not args.no_photon_extra:
            try:
                extra = photon_frame_extra(world, peak, cpu_baseline=not args.no_cpu_baseline)
            except Exception as e:                      # the headline line must survive a failure of the extra
                extra = {"error": "%s: %s" % (type(e).__name__, e)}
        per_launch_ms = ms_closest / args.steps
        achieved = 48.0 * n1 / (per_launch_ms * 1e-3) / 1e9
        value = (N1 + N2) * args.steps / (ms_total_m * 1e-3) / 1e6
        line = {
            "metric": METRIC, "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total_m / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, len(idx)),
            "closest_hit_mrays": N1 * args.steps / (ms_closest_m * 1e-3) / 1e6,
            "any_hit_mrays": N2 * args.steps / (ms_any_m * 1e-3) / 1e6,
            "incoherent_closest_hit_mrays": inc_mrays,
            "rays_per_step": {"closest": int(N1), "any": int(N2)},
            "accel": {"n_prims": info.n_prims, "wide_nodes": info.n_wide_nodes, "build_ms": info.build_ms, "upload_ms": info.upload_ms,
                      "sah_cost": info.sah_cost},
            "e2e": {"value": (N1 + N2) * e2e_steps / (e2e_ms_m * 1e-3) / 1e6, "unit": "Mrays/s",
                    "h2d_bytes_per_step": int(32 * (n1 + n2)), "d2h_bytes_per_step": int(16 * n1 + n2),
                    "ms_per_step": e2e_ms_m / e2e_steps, "api": "b2rt_trace_closest/b2rt_trace_any(B2RT_HOST)"},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": TRACE_KERNEL, "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": TRAFFIC_BYTES if (args.tris, n1) == (1_000_000, 2073600) else None,
                         "traffic_source": TRAFFIC_SOURCE, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": 48 * n1,
                         "note": "traversal is latency/issue bound, not HBM bound (SURVEY 8(d)); see profiles/"},
            "clocks": clocks, "wall_s_timed_region": wall,
        }
        if extra:
            line["photon_frame"] = extra
        if not args.no_cpu_baseline:
            smp = args.cpu_sample or None
            cb = cpu_reference_step(P, idx, rays, shadow, None, sample=smp)
            ref_hits = cb.pop("_hits", None)
            line["cpu_baseline"] = cb
            if ref_hits is not None:
                # self-certifying headline: the hits of the timed step against the reference's answers for the same rays
                line["parity"] = hit_parity(gpu_hits_np, gpu_occ_np, ref_hits)
                ref_hits["close"]()
        emit(line)


if __name__ == "__main__":
    main()
