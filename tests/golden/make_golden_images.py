"""Golden images from the reference's own CPU renderer (oracle/_ref/pbrt, built from /root/reference): the Cornell box
with two matte boxes, glass and mirror spheres, disk area light -- `directlighting` and `photonmap` (finalgather off).
Run in the build container:  python tests/golden/make_golden_images.py   -> tests/golden/images_cornell.npz (64x64)."""
import os
import subprocess
import sys
import tempfile
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import scenes as S         # noqa: E402

RES, SPP, NC, NI, NUSED, MAXD = 64, 1024, 20000, 300000, 100, 0.4


def main():
    desc = S.cornell_scene(boxes=True)
    out = {}
    with tempfile.TemporaryDirectory() as td:
        for name, integ in (("direct", 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]'),
                            ("photon", 'SurfaceIntegrator "photonmap" "integer causticphotons" [%d] "integer indirectphotons" [%d] '
                                       '"integer nused" [%d] "float maxdist" [%g] "bool finalgather" ["false"]' % (NC, NI, NUSED, MAXD))):
            pf = os.path.join(td, name + ".pfm")
            sc = os.path.join(td, name + ".pbrt")
            open(sc, "w").write(S.to_pbrt(desc, (0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, RES, RES, SPP, integ, pf))
            subprocess.run([os.path.join(ROOT, "oracle/_ref/pbrt"), "--quiet", sc], check=True, capture_output=True)
            out[name] = S.read_pfm(pf)
            print(name, out[name].mean((0, 1)))
    np.savez_compressed(os.path.join(ROOT, "tests/golden/images_cornell.npz"), params=np.array([SPP, NC, NI, NUSED]),
                        maxdist=np.float32(MAXD), **out)


if __name__ == "__main__":
    main()
