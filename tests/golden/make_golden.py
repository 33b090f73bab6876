"""Generates tests/golden/*.npz by running the REFERENCE's own CPU code (pbrt-v2 compiled from
/root/reference into oracle/_ref/ by `make -C oracle ref`) on seeded synthetic inputs.

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The inputs are regenerated from seeds by tests (cuda_raytrace_b200.scenes), so only rays + outputs
are stored.  Stored per case: rays, pbrt BVHAccel (t, prim, uv), pbrt exhaustive-loop (t, prim),
IntersectP occlusion, tie count.
"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def xf_translate(tx, ty, tz):
    o2w = np.array([1, 0, 0, tx, 0, 1, 0, ty, 0, 0, 1, tz], np.float32)
    w2o = np.array([1, 0, 0, -tx, 0, 1, 0, -ty, 0, 0, 1, -tz], np.float32)
    return o2w, w2o


def xf_disk_down(y):
    """pbrt `Translate 0 y 0  Rotate 90 1 0 0`: object z axis -> world -y."""
    o2w = np.array([1, 0, 0, 0, 0, 0, -1, y, 0, 1, 0, 0], np.float32)
    w2o = np.array([1, 0, 0, 0, 0, 0, 1, 0, 0, -1, 0, y], np.float32)
    return o2w, w2o


def case_scene(name):
    """Returns a list of ('mesh', P, idx) / ('sphere', o2w, w2o, r) / ('disk', o2w, w2o, h, r, inner)."""
    if name == "terrain_small":
        P, idx = S.heightfield_mesh(40, 40, None, seed=5)
        return [("mesh", P, idx), ("sphere", *xf_translate(0.2, 0.5, 0.1), 0.3),
                ("disk", *xf_disk_down(0.9), 0.0, 0.4, 0.0)]
    if name == "terrain_100k":
        P, idx = S.cfg2_mesh(100_000, seed=1234)
        return [("mesh", P, idx)]
    if name == "cornell":
        items = []
        for P, idx, _ in S.cornell_geometry():
            items.append(("mesh", P, idx))
        items.append(("sphere", *xf_translate(0.4, 0.35, 0.3), 0.35))
        items.append(("sphere", *xf_translate(-0.45, 0.35, -0.3), 0.35))
        items.append(("disk", *xf_disk_down(1.98), 0.0, 0.3, 0.0))
        return items
    raise KeyError(name)


def case_rays(name):
    if name == "terrain_small":
        return np.concatenate([S.pinhole_rays(64, 48), S.incoherent_rays(3000, [-1.2, -0.5, -1.2], [1.2, 1.2, 1.2])])
    if name == "terrain_100k":
        return np.concatenate([S.pinhole_rays(160, 90), S.incoherent_rays(20000, [-1.0, -0.3, -1.0], [1.0, 0.5, 1.0])])
    if name == "cornell":
        return np.concatenate([S.pinhole_rays(96, 96, eye=(0, 1, 3.4), look=(0, 1, 0), fov=40),
                               S.incoherent_rays(8000, [-0.99, 0.01, -0.99], [0.99, 1.97, 0.99], seed=3)])
    raise KeyError(name)


def fill(scene, items):
    for it in items:
        if it[0] == "mesh":
            scene.add_mesh(it[1], it[2])
        elif it[0] == "sphere":
            scene.add_sphere(it[1], it[2], it[3])
        else:
            scene.add_disk(it[1], it[2], it[3], it[4], it[5])


def main():
    for name in ("terrain_small", "terrain_100k", "cornell"):
        items = case_scene(name)
        rays = case_rays(name)
        rs = O.RefScene()
        fill(rs, items)
        rs.build()
        t, prim, uv, _ = rs.intersect(rays, 1)
        occ, _ = rs.occluded(rays, 1)
        brute_n = len(rays) if name != "terrain_100k" else 4000   # exhaustive loop is O(rays*prims)
        sel = np.arange(len(rays)) if brute_n == len(rays) else \
            np.random.default_rng(0).choice(len(rays), brute_n, replace=False)
        sel.sort()
        tb, pb, nties = rs.intersect_brute(rays[sel], 8)
        np.savez_compressed(os.path.join(OUT, name + ".npz"), rays=rays, bvh_t=t, bvh_prim=prim, bvh_uv=uv,
                            occluded=occ, brute_sel=sel.astype(np.int32), brute_t=tb, brute_prim=pb,
                            nties=np.int64(nties))
        print(name, "rays", len(rays), "hits", int((prim >= 0).sum()), "bvh!=brute",
              int((prim[sel] != pb).sum()), "ties", nties)


if __name__ == "__main__":
    main()
