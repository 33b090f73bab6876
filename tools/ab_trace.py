"""A/B timing of the traversal kernel variants on the Cfg2 rays (device-resident, L2 flushed between launches).
Variant comes from the environment (B2RT_TRACE_VARIANT, B2RT_TRACE_REFILL); prints one line.
python tools/ab_trace.py [tris]"""
import os
import sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

tris = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
P, idx = S.cfg2_mesh(tris)
rays = S.pinhole_rays(1920, 1080)
ctx = b2rt.Context(0, seed=1)
ctx.add_mesh(P, idx)
info = ctx.build_accel()
hits0 = ctx.trace_closest(rays)
shadow = S.shadow_rays(rays, hits0["t"], hits0["prim"], (0.5, 3.0, 0.5))
inc = S.incoherent_rays(len(rays), info.bounds[0:3], info.bounds[3:6], seed=99)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
with torch.cuda.stream(stream):
    d = {k: torch.from_numpy(v).cuda() for k, v in (("primary", rays), ("shadow", shadow), ("incoherent", inc))}
    d_hits = torch.empty((len(rays), 4), dtype=torch.float32, device="cuda")
    d_occ = torch.empty(len(rays), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ctx.sync()
out = []
ref = {}
for name, anyhit in (("primary", False), ("shadow", True), ("incoherent", False)):
    r = d[name]
    n = r.shape[0]
    fn = (lambda: ctx.trace_any_device(n, r.data_ptr(), d_occ.data_ptr())) if anyhit else \
         (lambda: ctx.trace_closest_device(n, r.data_ptr(), d_hits.data_ptr()))
    for _ in range(3):
        fn()
    ts = []
    for _ in range(20):
        with torch.cuda.stream(stream):
            flush.fill_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream)
        ctx.sync()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    out.append("%s %.1f Mrays/s (%.3f ms)" % (name, n / ms / 1e3, ms))
    import hashlib
    buf = d_occ[:n] if anyhit else d_hits[:n]
    out.append("sha " + hashlib.sha256(buf.cpu().numpy().tobytes()).hexdigest()[:12])
print("variant=%s refill=%s tris=%d | %s" % (os.environ.get("B2RT_TRACE_VARIANT", "2 (default)"), os.environ.get("B2RT_TRACE_REFILL", "0"),
                                            tris, " | ".join(out)), flush=True)
ctx.close()
