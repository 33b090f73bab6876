"""Object instances (two-level scene): the layered oracle (numpy restatement of pbrt's Transform(Ray) + the exhaustive C
loop in the instance's space) against the REFERENCE's own ObjectInstance path -- TransformedPrimitive over the object's
BVHAccel (core/primitive.cpp:95-120) -- from the committed golden (tests/golden/make_golden_instances.py) and, where
oracle/_ref exists, live."""
import numpy as np
import pytest

from oracle import oracle as O
from cuda_raytrace_b200 import scenes as S
import make_golden_instances as G
from helpers import load_golden


def _port(desc):
    ps = O.PortInstancedScene()
    S.fill_instanced(ps, desc, style="port")
    return ps


def _check(ps, r, t, prim, inst, occ):
    pt, pp, _, _, pi = ps.intersect(r, 4)
    bases = np.asarray(ps.instance_bases(), np.int32)
    local = np.where(pi >= 0, pp - bases[np.maximum(pi, 0)], pp)
    # pbrt's BVHAccel tests node boxes without padding and can lose a hit that the exhaustive loop accepts; the oracle is the
    # exhaustive loop.  Everything else must agree bit for bit.
    same = pt.view(np.uint32) == t.view(np.uint32)
    assert (~same).sum() <= 2, int((~same).sum())
    assert np.all(pt[~same] < t[~same])                     # only ever a nearer hit that BVHAccel dropped
    ok = same & (pp >= 0)
    tie = ok & ((pi != inst) | (local != prim))
    assert tie.sum() <= 3, int(tie.sum())                   # exact-t ties between coincident primitives
    assert (inst >= 0).sum() > 1000 and ((inst < 0) & (prim >= 0)).sum() > 1000
    assert np.array_equal(ps.occluded(r, 4)[same], occ[same])


def test_port_instances_match_pbrt_golden(built):
    g = load_golden("instances")
    _check(_port(G.scene()), G.rays(), g["t"], g["prim"], g["inst"], g["occ"])


def test_port_instances_match_pbrt_live(built):
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    desc = S.instanced_scene(n_inst=7, seed=11)
    r = S.incoherent_rays(4000, [-4, -1, -4], [4, 2, 4], seed=3)
    rs = O.RefScene()
    S.fill_instanced(rs, desc, style="ref")
    rs.build()
    t, prim, inst = rs.intersect_inst(r, 2)
    occ, _ = rs.occluded(r, 2)
    rs.close()
    ps = _port(desc)
    pt, pp, _, _, pi = ps.intersect(r, 2)
    same = pt.view(np.uint32) == t.view(np.uint32)
    assert (~same).sum() <= 2 and np.array_equal(pi[same], inst[same])


def test_exact_instances_equal_their_flattened_copy(built):
    """integer translations of grid-quantised meshes: instance space and world space give the same bits"""
    desc = S.instanced_scene(n_inst=9, seed=4, exact=True)
    r = S.pinhole_rays(40, 30, eye=(0.5, 3.0, 8.0), look=(0.0, 0.0, 0.0), fov=50)
    r[:, :3] = np.float32([0.5, 3.0, 8.0])
    ps = _port(desc)
    t, prim, b1, b2, inst = ps.intersect(r, 2)
    flat = O.PortScene()
    for sh in S.flatten_instanced(desc):
        flat.add_mesh(sh[1], sh[2])
    ft, fp, fb1, fb2 = flat.intersect(r, 2)
    ni = len(desc["instances"])
    assert np.array_equal(ft.view(np.uint32), t.view(np.uint32)) and np.array_equal(fb1, b1) and np.array_equal(fb2, b2)
    assert np.array_equal(np.where(inst >= 0, prim - ni, prim), fp)
    assert (inst >= 0).sum() > 100
