"""Exchange step of the sharded photon pass per mode (B2RT_MGPU_EXCHANGE): host wall time of the commit + exchange phase and
the bus bandwidth it implies, Cfg3 scene.  python tools/bench_exchange.py <n_gpus> [n_paths]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n_paths = int(sys.argv[2]) if len(sys.argv) > 2 else 24_000_000
desc = S.cornell_scene(spheres=True, wide=1.78)
for mode in ("fused", "push", "copy", "nccl"):
    os.environ["B2RT_MGPU_EXCHANGE"] = mode
    with b2rt.MultiContext(G, seed=3) as m:
        m.fill(lambda c: S.fill_scene(c, desc))
        m.build_accel()
        rows = []
        for rep in range(5):
            m.photons_clear()
            m.stats(reset=True)
            nc, ni = m.photons_trace(0, n_paths)
            st = m.stats(reset=True)
            if rep:
                rows.append(st)
        ex = sorted(r["exchange_ms"] for r in rows)[len(rows) // 2]
        tr = sorted(r["trace_ms"] for r in rows)[len(rows) // 2]
        byts = rows[0]["exchange_bytes"]
        print(json.dumps({"gpus": G, "mode": m.info()["exchange"], "photons": nc + ni, "trace_ms": tr, "commit_exchange_ms": ex,
                          "bus_GBs": (G - 1) / G * byts / (ex * 1e-3) / 1e9 if G > 1 else None}), flush=True)
