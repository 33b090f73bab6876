"""Traversal work per ray from the -DB2RT_STATS build: wide nodes fetched and primitive records tested
by trace_spec_kernel, and the L2 -> SM bytes per ray they imply (128 B per node, 48 B per primitive record).
    B2RT_LIB=cuda_raytrace_b200/libb2rt_stats.so python tools/trace_stats.py [n_tris]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("B2RT_LIB", os.path.join(ROOT, "cuda_raytrace_b200", "libb2rt_stats.so"))
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

ntris = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
P, idx = S.cfg2_mesh(ntris)
rays = S.pinhole_rays(1920, 1080)
with b2rt.Context(0, seed=1) as ctx:
    ctx.add_mesh(P, idx)
    info = ctx.build_accel()
    hits = ctx.trace_closest(rays)
    ctx.trace_stats(reset=True)
    out = {"n_tris": ntris, "wide_nodes": info.n_wide_nodes}
    shadow = S.shadow_rays(rays, hits["t"], hits["prim"], (0.5, 3.0, 0.5))
    inc = S.incoherent_rays(len(rays), info.bounds[0:3], info.bounds[3:6], seed=99)
    for name, r, fn in (("primary_closest", rays, ctx.trace_closest), ("shadow_any", shadow, ctx.trace_any),
                        ("incoherent_closest", inc, ctx.trace_closest)):
        fn(r)
        nodes, prims = ctx.trace_stats(reset=True)
        out[name] = {"rays": len(r), "nodes_per_ray": nodes / len(r), "prims_per_ray": prims / len(r),
                     "l2_to_sm_bytes_per_ray": (nodes * 128 + prims * 48) / len(r)}
print(json.dumps(out))
