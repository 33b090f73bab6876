"""Two-level BVH against the flattened copy of the same scene: device memory, build time and closest-hit throughput.
Scene: NI instances (random rotation, scale, translation over a plane) of one NT-triangle terrain patch + a ground mesh.
python tools/bench_instances.py [n_instances] [tris_per_object] [width] [height]      -> one JSON line per variant"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

NI = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
NT = int(sys.argv[2]) if len(sys.argv) > 2 else 20000
W = int(sys.argv[3]) if len(sys.argv) > 3 else 1920
H = int(sys.argv[4]) if len(sys.argv) > 4 else 1080

side = int(np.sqrt(NT / 2))
Po, io = S.heightfield_mesh(side, side, None, seed=3)
Po = Po.copy(); Po[:, 1] = Po[:, 1] * 0.6
Pg, ig = S.heightfield_mesh(64, 64, None, seed=4)
Pg = Pg.copy(); Pg[:, [0, 2]] *= 40.0; Pg[:, 1] = Pg[:, 1] * 0.5 - 1.5
rng = np.random.default_rng(1)
inst = []
g = int(np.ceil(np.sqrt(NI)))
for k in range(NI):
    t = ((k % g) / g * 70.0 - 35.0 + rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 2.0), (k // g) / g * 70.0 - 35.0 + rng.uniform(-0.5, 0.5))
    inst.append(S.xf_rotate_scale(rng.normal(size=3), rng.uniform(0, 360), rng.uniform(0.6, 1.4, 3), t))
desc = {"world": [("mesh", Pg.astype(np.float32), ig, None, None)], "objects": [[("mesh", Po.astype(np.float32), io, None, None)]],
        "instances": [(0, a, b) for (a, b) in inst]}
rays = S.pinhole_rays(W, H, eye=(0.0, 30.0, 60.0), look=(0.0, 0.0, 0.0), fov=50)
inc = S.incoherent_rays(1 << 21, [-35, -2, -35], [35, 4, 35], seed=5)


def run(name, fill):
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    with b2rt.Context(0, seed=1) as c:
        fill(c)
        info = c.build_accel()
        c.sync()
        used = free0 - torch.cuda.mem_get_info()[0]
        out = {"variant": name, "instances": NI, "tris_per_object": len(io), "virtual_tris": NI * len(io) + len(ig), "tree_items": info.n_prims,
               "device_MB": round(used / 2 ** 20, 1), "build_ms": round(info.build_ms, 3), "upload_ms": round(info.upload_ms, 3)}
        for tag, r in (("camera", rays), ("incoherent", inc)):
            d = torch.from_numpy(r).cuda()
            h = torch.empty((len(r), 4), dtype=torch.float32, device="cuda")
            torch.cuda.synchronize()
            for _ in range(3):
                c.trace_closest_device(len(r), d.data_ptr(), h.data_ptr())
            ts = []
            for _ in range(7):
                c.timer_start()
                c.trace_closest_device(len(r), d.data_ptr(), h.data_ptr())
                ts.append(c.timer_stop())
            ms = float(np.median(ts))
            hh = h.cpu().numpy()
            out[tag + "_ms"] = round(ms, 3)
            out[tag + "_mrays"] = round(len(r) / ms / 1e3, 1)
            out[tag + "_hit_frac"] = round(float(np.isfinite(hh[:, 0]).mean()), 4)
            out[tag + "_t_sum"] = float(np.where(np.isfinite(hh[:, 0]), hh[:, 0], 0).astype(np.float64).sum())
    print(json.dumps(out), flush=True)
    return out


a = run("two_level", lambda c: S.fill_instanced(c, desc))
if NI * len(io) <= 120_000_000:
    def flat(c):
        for sh in S.flatten_instanced(desc):
            c.add_mesh(sh[1], sh[2])
    b = run("flattened", flat)
    print(json.dumps({"memory_ratio": round(b["device_MB"] / max(a["device_MB"], 1e-9), 1), "build_ratio": round(b["build_ms"] / a["build_ms"], 2),
                      "camera_speed_ratio_two_level_over_flat": round(a["camera_mrays"] / b["camera_mrays"], 3),
                      "incoherent_speed_ratio_two_level_over_flat": round(a["incoherent_mrays"] / b["incoherent_mrays"], 3),
                      "t_sum_rel_diff": abs(a["camera_t_sum"] - b["camera_t_sum"]) / max(abs(b["camera_t_sum"]), 1e-30)}))
