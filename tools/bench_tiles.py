"""Camera pass of the Cfg3 frame through b2rt_multi_render_frame for several image-tile edges (B2RT_MGPU_TILE): load balance
of the interleaved tiles against the coherence of the rays inside one.  Device times (CUDA events on every GPU's stream, max
over GPUs, median of 5); the frame's SHA-256 must not depend on the tile edge.
python tools/bench_tiles.py [n_gpus]"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else b2rt.load().b2rt_device_count()
desc = S.cornell_scene(spheres=True, wide=1.78)
cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 1920, 1080)
with b2rt.MultiContext(n, seed=3) as m:
    m.fill(lambda c: S.fill_scene(c, desc))
    m.build_accel()
    frame = m.ctxs[0].pinned_empty((1080, 1920, 3))
    m.photons_trace(0, 24_000_000)
    m.grid_build(0); m.grid_build(1)
    for tile, host in ((64, 1), (64, 0), (16, 1), (32, 1), (128, 1), (64, 1), (64, 0)):
        os.environ["B2RT_MGPU_TILE"] = str(tile)
        os.environ["B2RT_MGPU_HOSTFRAME"] = str(host)       # 1: the GPUs store their pixels straight into the pinned frame
        ts = []
        for rep in range(6):
            for c in m.ctxs:
                c.timer_start()
            m.render_frame(cam, spp=1, nused=64, max_dist=0.1, out=frame)
            per = [c.timer_stop() for c in m.ctxs]
            if rep:
                ts.append(per)
        per = np.median(np.asarray(ts), axis=0)
        print(json.dumps({"gpus": n, "tile": tile, "direct_host_frame": host, "camera_pass_ms": float(per.max()), "per_gpu_ms": [round(float(x), 3) for x in per],
                          "imbalance": float(per.max() / per.mean()), "sha256": hashlib.sha256(np.ascontiguousarray(frame).tobytes()).hexdigest()[:16]}),
              flush=True)
