"""Host-buffer path (e2e): pinned-copy bandwidth of the box and the Cfg2 step through b2rt_trace_*(B2RT_HOST) for the
chunk size in B2RT_HOST_CHUNK.  Prints one line."""
import os
import sys
import time
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

n = 64 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
bw = {}
for name, (dst, src) in (("h2d", (d, h)), ("d2h", (h, d))):
    for _ in range(2):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    bw[name] = 10 * n / (time.perf_counter() - t0) / 1e9
P, idx = S.cfg2_mesh(1_000_000)
rays = S.pinhole_rays(1920, 1080)
ctx = b2rt.Context(0, seed=1)
ctx.add_mesh(P, idx)
ctx.build_accel()
hits0 = ctx.trace_closest(rays)
shadow = S.shadow_rays(rays, hits0["t"], hits0["prim"], (0.5, 3.0, 0.5))
n1, n2 = len(rays), len(shadow)
h_rays, h_shadow = torch.from_numpy(rays).pin_memory(), torch.from_numpy(shadow).pin_memory()
h_hits, h_occ = torch.empty((n1, 4), dtype=torch.float32).pin_memory(), torch.empty(n2, dtype=torch.uint8).pin_memory()


def step():
    rc = ctx.lib.b2rt_trace_closest(ctx.h, n1, h_rays.data_ptr(), h_hits.data_ptr(), b2rt.HOST)
    rc |= ctx.lib.b2rt_trace_any(ctx.h, n2, h_shadow.data_ptr(), h_occ.data_ptr(), b2rt.HOST)
    assert rc == 0


for _ in range(3):
    step()
t0 = time.perf_counter()
for _ in range(50):
    step()
ms = (time.perf_counter() - t0) / 50 * 1e3
floor = (32 * (n1 + n2)) / (bw["h2d"] * 1e9) * 1e3
print("chunk=%s | pinned copy h2d %.1f GB/s d2h %.1f GB/s | step %.3f ms = %.0f Mrays/s | upload-only floor %.3f ms"
      % (os.environ.get("B2RT_HOST_CHUNK", "262144"), bw["h2d"], bw["d2h"], ms, (n1 + n2) / ms / 1e3, floor), flush=True)
ctx.close()
