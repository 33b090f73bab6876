"""Golden hit attributes (K6) from the REFERENCE's own CPU code: pbrt-v2 Triangle::Intersect + GetShadingGeometry +
BSDF::BSDF through oracle/_ref (ref_capi.cpp ref_surface), for triangle meshes with per-vertex normals and / or uv and
with reversed orientation.  Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_surface.py        -> tests/golden/surface.npz
Scenes and rays are regenerated from seeds by the tests; only pbrt's answers are stored."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "surface.npz")
# (name, normals, uv, reverse)
CASES = (("n_uv", True, True, False), ("n", True, False, False), ("uv", False, True, False),
         ("n_uv_rev", True, True, True), ("rev", False, False, True))


def mesh():
    """a smooth sphere mesh next to a bumpy heightfield with uv = its (x, z) parametrisation and area-weighted normals"""
    P, idx, N, uv = S.sphere_mesh(24, 12, 0.5, (0.1, 0.2, -0.1), seed=3)
    Ph, ih = S.heightfield_mesh(12, 12, None, seed=9)
    Ph = Ph.copy(); Ph[:, 1] -= 0.6
    uvh = (Ph[:, [0, 2]] * 0.5 + 0.5).astype(np.float32)
    Nh = np.zeros_like(Ph)
    for t in ih:
        fn = np.cross(Ph[t[2]] - Ph[t[0]], Ph[t[1]] - Ph[t[0]])
        Nh[t] += fn
    Nh /= np.maximum(np.linalg.norm(Nh, axis=1, keepdims=True), 1e-20)
    return (P, idx, N, uv), (Ph, ih, Nh.astype(np.float32), uvh)


def rays():
    return np.concatenate([S.incoherent_rays(6000, [-1, -1, -1], [1, 1, 1], seed=5),
                           S.pinhole_rays(48, 48, eye=(0.1, 0.4, 2.0), look=(0.1, 0.0, -0.1), fov=45)])


def fill(target, case):
    """target: anything with add_mesh(P, idx, N=, uv=, reverse=) -- b2rt.Context, oracle.PortFullScene"""
    _, n, u, rev = case
    for (P, idx, N, uv) in mesh():
        target.add_mesh(P, idx, N=N if n else None, uv=uv if u else None, reverse=rev)


def main():
    out = {}
    r = rays()
    for case in CASES:
        rs = O.RefScene()
        _, n, u, rev = case
        for (P, idx, N, uv) in mesh():
            rs.add_mesh_attr(P, idx, N if n else None, uv if u else None, rev)
        rs.build()
        out[case[0]] = rs.surface_at(r, 4)
        rs.close()
        print(case[0], "hits", int((out[case[0]]["prim"] >= 0).sum()), "of", len(r))
    np.savez_compressed(OUT, **out)


if __name__ == "__main__":
    main()
