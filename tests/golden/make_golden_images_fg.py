"""Golden image from the reference's own CPU renderer (oracle/_ref/pbrt, built from /root/reference) for pbrt's DEFAULT
photonmap mode, final gathering on: the Cornell box of make_golden_images.py, `photonmap` with finalgather true,
finalgathersamples 32, gatherangle 10.
Run in the build container:  python tests/golden/make_golden_images_fg.py   -> tests/golden/images_cornell_fg.npz (64x64)."""
import os
import subprocess
import sys
import tempfile
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from cuda_raytrace_b200 import scenes as S         # noqa: E402

RES, SPP, NC, NI, NUSED, MAXD, GS, GA = 64, 1024, 20000, 300000, 100, 0.4, 32, 10.0


def main():
    desc = S.cornell_scene(boxes=True)
    integ = ('SurfaceIntegrator "photonmap" "integer causticphotons" [%d] "integer indirectphotons" [%d] "integer nused" [%d] '
             '"float maxdist" [%g] "bool finalgather" ["true"] "integer finalgathersamples" [%d] "float gatherangle" [%g]'
             % (NC, NI, NUSED, MAXD, GS, GA))
    with tempfile.TemporaryDirectory() as td:
        pf, sc = os.path.join(td, "fg.pfm"), os.path.join(td, "fg.pbrt")
        open(sc, "w").write(S.to_pbrt(desc, (0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, RES, RES, SPP, integ, pf))
        subprocess.run([os.path.join(ROOT, "oracle/_ref/pbrt"), "--quiet", sc], check=True, capture_output=True)
        img = S.read_pfm(pf)
    print("fg", img.mean((0, 1)))
    np.savez_compressed(os.path.join(ROOT, "tests/golden/images_cornell_fg.npz"),
                        params=np.array([SPP, NC, NI, NUSED, GS]), maxdist=np.float32(MAXD), gatherangle=np.float32(GA), fg=img)


if __name__ == "__main__":
    main()
