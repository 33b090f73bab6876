"""Renders the Cornell scene with b2rt and with the reference's CPU renderer (oracle/_ref/pbrt) and prints image
metrics.  Test/diagnostic tool (uses oracle/_ref): python tools/compare_render.py [res] [spp]"""
import os
import subprocess
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import b2rt, scenes as S   # noqa: E402

res = int(sys.argv[1]) if len(sys.argv) > 1 else 128
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 64
out = os.path.join(ROOT, "gpurun_out")
os.makedirs(out, exist_ok=True)
EYE, LOOK, UP, FOV = (0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40
NC, NI, NUSED, MAXD = 20000, 200000, 50, 0.3
desc = S.cornell_scene()
cases = {
    "direct": 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]',
    "photon": 'SurfaceIntegrator "photonmap" "integer causticphotons" [%d] "integer indirectphotons" [%d] "integer nused" [%d] '
              '"float maxdist" [%g] "bool finalgather" ["false"] "integer maxphotondepth" [5] "integer maxspeculardepth" [5]'
              % (NC, NI, NUSED, MAXD),
}
ref = {}
for name, integ in cases.items():
    pf = os.path.join(out, "ref_%s.pfm" % name)
    open(os.path.join(out, "ref_%s.pbrt" % name), "w").write(S.to_pbrt(desc, EYE, LOOK, UP, FOV, res, res, spp, integ, pf))
    t = time.time()
    subprocess.run([os.path.join(ROOT, "oracle/_ref/pbrt"), "--quiet", os.path.join(out, "ref_%s.pbrt" % name)],
                   check=True, capture_output=True)
    ref[name] = S.read_pfm(pf)
    print("pbrt %s: %.2f s" % (name, time.time() - t))

with b2rt.Context(0, seed=3) as ctx:
    S.fill_scene(ctx, desc)
    ctx.build_accel()
    cam = b2rt.pbrt_camera(EYE, LOOK, UP, FOV, res, res)
    t = time.time()
    img_d = ctx.render_frame(cam, spp=spp, caustic=False, indirect=False)
    print("b2rt direct: %.3f s" % (time.time() - t), S.image_metrics(img_d, ref["direct"]))
    t = time.time()
    (pc, pi), (nc, ni) = ctx.photons_shoot(NC, NI, batch_paths=65536)
    g0 = ctx.grid_build(0)
    g1 = ctx.grid_build(1)
    print("shoot+grid %.3f s  paths %d %d stored %d %d cells %s %s" % (time.time() - t, pc, pi, nc, ni, list(g0.dims), list(g1.dims)))
    t = time.time()
    img_p = ctx.render_frame(cam, spp=spp, nused=NUSED, max_dist=MAXD)
    print("b2rt photon: %.3f s" % (time.time() - t), S.image_metrics(img_p, ref["photon"]))
    for nm, a, b in (("direct", img_d, ref["direct"]), ("photon", img_p, ref["photon"])):
        print(nm, "mean b2rt", a.mean((0, 1)), "pbrt", b.mean((0, 1)))
        h = res // 4
        for qy in range(4):
            print("  ", " ".join("%.3f/%.3f" % (a[qy*h:(qy+1)*h, qx*h:(qx+1)*h].mean(), b[qy*h:(qy+1)*h, qx*h:(qx+1)*h].mean()) for qx in range(4)))
    np.savez_compressed(os.path.join(out, "render_cmp.npz"), img_d=img_d, img_p=img_p, ref_d=ref["direct"], ref_p=ref["photon"])
