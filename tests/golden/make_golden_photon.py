"""Generates tests/golden/photon_cornell.npz by running the REFERENCE's own CPU photon mapper (pbrt-v2
PhotonIntegrator, oracle/ref_photon.cpp over oracle/_ref/libpbrt_oracle.so) in the build container.

  * pbrt's own shooting (Halton + MT19937): flux per emitted path and stored photons per path for the caustic and
    indirect maps -- the statistical pin for our Philox-driven restatement / GPU kernel.
  * pbrt's own estimator (KdTree<Photon>::Lookup + LPhoton) evaluated on a photon set produced by the C oracle
    (seed 5, 100 000 paths) at seeded query points -- the value pin for port_gather / the GPU gather kernel.
Run: python tests/golden/make_golden_photon.py      (needs /root/reference to have been built into oracle/_ref)
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SEED, NPATHS = 5, 100_000
GATHER_CASES = [(50, 0.09), (64, 0.01), (20, 1e-4), (0, 2.5e-3)]     # (k, max_dist2); k = 0: fixed radius


def queries(nq=400, seed=0):
    rng = np.random.default_rng(seed)
    h = nq // 2
    q = np.zeros((nq, 12), np.float32)
    q[:h, 0], q[:h, 2] = rng.uniform(-.95, .95, h), rng.uniform(-.95, .95, h)
    q[:h, 3:6] = (0, 1, 0)                                        # floor
    q[h:, 0], q[h:, 1], q[h:, 2] = -1, rng.uniform(.05, 1.95, nq - h), rng.uniform(-.95, .95, nq - h)
    q[h:, 3:6] = (1, 0, 0)                                        # left wall
    q[:, 6:9] = (0.7, 0.6, 0.5)
    return q


def oracle_photons():
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    return sc.photon_trace(SEED, 0, NPATHS)


def main():
    rs = O.RefPhotonScene()
    S.fill_scene(rs, S.cornell_scene())
    rc, ri, (pc, pi) = rs.shoot(60000, 1000000, ncores=1)
    out = {"pbrt_caustic_flux_per_path": rc[:, 3:6].sum(0) / pc, "pbrt_indirect_flux_per_path": ri[:, 3:6].sum(0) / pi,
           "pbrt_caustic_per_path": np.float64(len(rc) / pc), "pbrt_indirect_per_path": np.float64(len(ri) / pi),
           "pbrt_n": np.array([len(rc), len(ri), pc, pi], np.int64)}
    _, ind = oracle_photons()
    g = O.grid_build(ind)
    q = queries()
    for n, (k, md2) in enumerate(GATHER_CASES):
        out["lphoton_%d" % n] = rs.lphoton(O.photons_to9(g["sorted"]), NPATHS, q, k if k > 0 else len(ind), md2)
    np.savez_compressed(os.path.join(OUT, "photon_cornell.npz"), **out)
    for k_, v in out.items():
        print(k_, v if np.size(v) < 8 else v.shape)


if __name__ == "__main__":
    main()
