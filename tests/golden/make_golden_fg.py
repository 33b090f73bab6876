"""Generates tests/golden/fg.npz by running the REFERENCE's own final-gathering pieces (pbrt-v2 photonmap.cpp compiled
unmodified into oracle/_ref, hooks in oracle/ref_photon.cpp) in the build container:

  * EPhoton (KdTree<Photon>::Lookup + the irradiance sum) on the C oracle's indirect photons (seed 5, 100 000 paths) at
    seeded query points -- the value pin for port_irradiance / the GPU irradiance kernel;
  * KdTree<RadiancePhoton>::Lookup + RadiancePhotonProcess with an unbounded radius on the C oracle's radiance photons,
    each photon's Lo.x set to its index -- the exact pin (which photon is found) for port_radiance_lookup / the GPU shell
    search.
Run: python tests/golden/make_golden_fg.py      (needs /root/reference to have been built into oracle/_ref)
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402
import make_golden_photon as G                     # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
E_CASES = [(50, 0.09), (50, 0.01), (8, 1e-3)]        # (nLookup, max_dist2)
FG_PATHS = 20_000


def fg_photons():
    """(direct, radiance) photons of the C oracle, paths [0, FG_PATHS) of seed G.SEED."""
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    return sc.photon_trace_fg(G.SEED, 0, FG_PATHS)


def lookup_queries(nq=600, seed=11):
    """Points on and off the Cornell box's surfaces with assorted normals (some far outside the photons' bounding box)."""
    rng = np.random.default_rng(seed)
    q = np.zeros((nq, 12), np.float32)
    q[:, 0:3] = rng.uniform(-1.0, 1.0, (nq, 3)) * (1.0, 1.0, 1.0) + (0.0, 1.0, 0.0)
    far = rng.random(nq) < 0.1
    q[far, 0:3] *= 4.0
    n = rng.normal(size=(nq, 3))
    axis = rng.random(nq) < 0.5
    n[axis] = np.eye(3)[rng.integers(0, 3, axis.sum())] * rng.choice([-1.0, 1.0], (axis.sum(), 1))
    q[:, 3:6] = n / np.linalg.norm(n, axis=1, keepdims=True)
    return q


def radiance9(rad):
    """{p, decoded normal, (index + 1, 0, 0)} rows for the reference's lookup."""
    out = np.zeros((len(rad), 9), np.float32)
    out[:, 0:3] = rad["p"]
    out[:, 3:6] = O.oct_decode(rad["wi_oct"])
    out[:, 6] = np.arange(len(rad)) + 1            # 0 = "no photon found"
    return out


def main():
    rs = O.RefPhotonScene()
    _, ind = G.oracle_photons()
    g = O.grid_build(ind)
    q = G.queries()
    out = {}
    for n, (k, md2) in enumerate(E_CASES):
        out["ephoton_%d" % n] = rs.ephoton(O.photons_to9(g["sorted"]), G.NPATHS, q, k, md2)
    _, rad = fg_photons()
    lq = lookup_queries()
    out["lookup_index"] = rs.radiance_lookup(radiance9(rad), lq)[:, 0].astype(np.int32) - 1
    out["n_radiance"] = np.int64(len(rad))
    np.savez_compressed(os.path.join(OUT, "fg.npz"), **out)
    for k_, v in out.items():
        print(k_, v if np.size(v) < 8 else v.shape)


if __name__ == "__main__":
    main()
