"""RMSE of the direct-lighting image (Cornell box, disk area light) against a converged image, per pixel pattern and spp:
what pbrt's stratified / low-discrepancy samplers buy over independent uniforms on this backend.  Prints one JSON line.
python tools/sampler_convergence.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

with b2rt.Context(0, seed=5) as c:
    S.fill_scene(c, S.cornell_scene(spheres=True))
    c.build_accel()
    cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 192, 128)
    kw = dict(caustic=False, indirect=False)
    c.set_sampler("lowdiscrepancy")
    ref = c.render_frame(cam, spp=8192, **kw).astype(np.float64)
    out = {"scene": "Cornell box + spheres, direct lighting only, 192x128", "reference": "lowdiscrepancy, 8192 spp",
           "rmse_relative_to_mean": {}}
    for name in ("random", "stratified", "lowdiscrepancy"):
        row = {}
        for spp in (1, 4, 16, 64):
            n = int(round(spp ** 0.5))
            c.set_sampler(name, n, n)
            img = c.render_frame(cam, spp=spp, **kw).astype(np.float64)
            row[str(spp)] = round(float(np.sqrt(((img - ref) ** 2).mean()) / ref.mean()), 4)
        out["rmse_relative_to_mean"][name] = row
print(json.dumps(out))
