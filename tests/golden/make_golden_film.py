"""Golden film images from the reference's own ImageFilm + pixel filters (oracle/_ref, built from /root/reference):
the seeded sample sets of film_cases.py fed through ImageFilm::AddSample / WriteImage (.pfm).
Run in the build container:  python tests/golden/make_golden_film.py  -> tests/golden/film.npz"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from oracle import oracle as O     # noqa: E402
import film_cases as F             # noqa: E402

out = {}
for n, (name, kw, xres, yres, spp, window) in enumerate(F.CASES):
    flt = O.filter_args(name, **kw)
    cells, xy, rgb, px = F.make_samples(flt, xres, yres, spp, window, seed=100 + n)
    out["film_%d" % n] = O.ref_film(flt, xres, yres, xy, rgb, px)
np.savez_compressed(os.path.join(ROOT, "tests/golden/film.npz"), **out)
print({k: v.shape for k, v in out.items()})
