"""Golden hits of an INSTANCED scene from the REFERENCE's own CPU code: pbrt-v2 ObjectInstance -> TransformedPrimitive over
the object's BVHAccel (core/api.cpp:1140-1172, core/primitive.cpp:95-120) through oracle/_ref (ref_capi.cpp
ref_scene_add_instance / ref_intersect_inst / ref_surface_inst).  Run in the build container only (needs /root/reference):
    python tests/golden/make_golden_instances.py        -> tests/golden/instances.npz
The scene and the rays are regenerated from seeds by the tests; only pbrt's answers are stored."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "instances.npz")


def scene():
    return S.instanced_scene(n_inst=10, seed=5)


def rays():
    return np.concatenate([S.incoherent_rays(8000, [-4, -1, -4], [4, 2, 4], seed=21),
                           S.pinhole_rays(64, 48, eye=(0.5, 3.0, 7.0), look=(0.0, 0.0, 0.0), fov=50)])


def main():
    desc, r = scene(), rays()
    rs = O.RefScene()
    S.fill_instanced(rs, desc, style="ref")
    rs.build()
    t, prim, inst = rs.intersect_inst(r, 4)
    occ, _ = rs.occluded(r, 4)
    surf, sinst = rs.surface_at_inst(r, 4)
    assert np.array_equal(sinst, inst)
    rs.close()
    print("hits", int((prim >= 0).sum()), "of", len(r), "instanced", int((inst >= 0).sum()))
    np.savez_compressed(OUT, t=t, prim=prim, inst=inst, occ=occ, surf=surf)


if __name__ == "__main__":
    main()
