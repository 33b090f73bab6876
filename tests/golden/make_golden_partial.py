"""Golden hits of PARTIAL spheres ("zmin" / "zmax") from the REFERENCE's own CPU code: pbrt-v2 Sphere::Intersect /
IntersectP (shapes/sphere.cpp:58-106, 159-205) under BVHAccel, through oracle/_ref.  Build container only:
    python tests/golden/make_golden_partial.py        -> tests/golden/partial_spheres.npz"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "partial_spheres.npz")


def shapes():
    """(o2w, w2o, radius, zmin, zmax, phimax): a bowl, a band cut to 250 degrees, a tilted and stretched cap, one full sphere,
    an orange slice (phimax 100 only); plus a ground mesh and a 3/4 disk with a hole"""
    Pg, ig = S.heightfield_mesh(10, 10, None, seed=2)
    Pg = Pg.copy(); Pg[:, 1] = Pg[:, 1] * 0.2 - 0.9; Pg[:, [0, 2]] *= 2.0
    sph = [S.xf_translate(-0.9, 0.0, 0.2) + (0.5, None, 0.1, None),
           S.xf_rotate_scale((1, 0, 0), -90.0, 1.0, (0.2, 0.1, -0.3)) + (0.45, -0.2, 0.25, 250.0),
           S.xf_rotate_scale((1, 1, 0.3), 40.0, (1.0, 0.7, 1.3), (1.0, 0.2, 0.4)) + (0.4, 0.1, None, None),
           S.xf_translate(0.1, 0.9, 0.6) + (0.3, -5.0, 5.0, None),
           S.xf_rotate_scale((0.2, 1, 0.1), 70.0, 1.0, (-0.3, 0.6, -0.8)) + (0.4, None, None, 100.0)]
    return (Pg.astype(np.float32), ig), sph


def disk():
    """(o2w, w2o, height, radius, inner, phimax)"""
    return S.xf_rotate_scale((1, 0.2, 0), -70.0, 1.0, (0.9, -0.2, -0.6)) + (0.05, 0.6, 0.15, 270.0)


def fill(target, style):
    (Pg, ig), sph = shapes()
    target.add_mesh(Pg, ig)
    for (o2w, w2o, r, z0, z1, pm) in sph:
        if style == "b2rt":
            target.add_sphere(o2w, r, w2o=w2o, zmin=z0, zmax=z1, phimax=pm)
        else:
            target.add_sphere(o2w, w2o, r, zmin=z0, zmax=z1, phimax=pm)
    o2w, w2o, h, r, ri, pm = disk()
    if style == "b2rt":
        target.add_disk(o2w, h, r, ri, w2o=w2o, phimax=pm)
    else:
        target.add_disk(o2w, w2o, h, r, ri, phimax=pm)


def rays():
    return np.concatenate([S.incoherent_rays(12000, [-1.8, -0.8, -1.2], [1.8, 1.4, 1.4], seed=8),
                           S.pinhole_rays(64, 48, eye=(0.1, 1.6, 3.2), look=(0.0, 0.0, 0.0), fov=50)])


def main():
    r = rays()
    rs = O.RefScene()
    fill(rs, "ref")
    rs.build()
    t, prim, _, _ = rs.intersect(r, 4)
    occ, _ = rs.occluded(r, 4)
    rs.close()
    print("hits", int((prim >= 0).sum()), "sphere / disk hits", int((prim >= prim.max() - 5).sum()))
    np.savez_compressed(OUT, t=t, prim=prim, occ=occ)


if __name__ == "__main__":
    main()
