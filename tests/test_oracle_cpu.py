"""CPU-only checks (-m "not gpu"): the oracle restatement against the golden vectors produced by the
reference's own pbrt-v2 code (tests/golden/make_golden.py), the restated builder's invariants, and the
C-ABI library's exported symbols.  No compute call into libb2rt.so happens here."""
import ctypes
import os
import re
import numpy as np
import pytest

from helpers import CASES, load_golden, case_scene, fill, bits
from oracle import oracle as O
from cuda_raytrace_b200 import scenes as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("name", CASES)
def test_port_matches_reference_golden(name, built):
    """oracle_port.c (min t, min id exhaustive loop) == pbrt's Shape::Intersect exhaustive loop, bit for bit,
    and == pbrt BVHAccel::Intersect except on exact-t ties (where pbrt's answer depends on traversal order)."""
    g = load_golden(name)
    ps = O.PortScene()
    fill(ps, case_scene(name))
    sel = g["brute_sel"]
    rays = g["rays"]
    t, prim, b1, b2 = ps.intersect(rays[sel], nthreads=8)
    assert np.array_equal(prim, g["brute_prim"])
    assert np.array_equal(bits(t), bits(g["brute_t"]))
    # vs pbrt's BVH: t identical everywhere; prim identical except for ties
    assert np.array_equal(bits(t), bits(g["bvh_t"][sel]))
    diff = int((prim != g["bvh_prim"][sel]).sum())
    assert diff <= int(g["nties"])
    # barycentrics vs pbrt's (u,v) with its default uvs (0,0),(1,0),(1,1): u = b1+b2, v = b2
    ntri = len(ps.idx)
    tri = (prim >= 0) & (prim < ntri) & (prim == g["bvh_prim"][sel])
    # (value equality: pbrt's 0*b0 + 0*b1 + b2 turns a -0.0 into +0.0)
    assert np.array_equal((b1 + b2)[tri], g["bvh_uv"][sel][tri, 0])
    assert np.array_equal(b2[tri], g["bvh_uv"][sel][tri, 1])


@pytest.mark.parametrize("name", ("terrain_small", "cornell"))
def test_port_occlusion_matches_reference_golden(name, built):
    g = load_golden(name)
    ps = O.PortScene()
    fill(ps, case_scene(name))
    occ = ps.occluded(g["rays"], nthreads=8)
    assert np.array_equal(occ, g["occluded"])


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref not built (no /root/reference here)")
def test_port_matches_live_reference(built):
    """Where the reference binary exists: fresh random inputs through both."""
    P, idx = S.heightfield_mesh(17, 23, None, seed=77)
    rays = S.incoherent_rays(2000, [-1, -0.3, -1], [1, 0.5, 1], seed=5)
    ps, rs = O.PortScene(), O.RefScene()
    ps.add_mesh(P, idx)
    rs.add_mesh(P, idx)
    rs.build()
    t, prim, _, _ = ps.intersect(rays, 4)
    tb, pb, _ = rs.intersect_brute(rays, 4)
    assert np.array_equal(prim, pb) and np.array_equal(bits(t), bits(tb))


def test_sort_restatement(built):
    rng = np.random.default_rng(0)
    for n, kb in ((0, 32), (1, 32), (1000, 8), (70000, 30), (5000, 32)):
        keys = rng.integers(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32)
        vals = np.arange(n, dtype=np.uint32)
        k, v = O.sort_pairs(keys, vals, kb)
        eff = 8 * ((kb + 7) // 8)      # whole 8-bit passes: bits above key_bits must be zero or are sorted too
        mask = np.uint32((1 << eff) - 1) if eff < 32 else np.uint32(0xFFFFFFFF)
        order = np.argsort(keys & mask, kind="stable")
        assert np.array_equal(v, order.astype(np.uint32))
        assert np.array_equal(k, keys[order])


def _check_tree(out, n):
    parent, child = out["bin_parent"], out["bin_child"].reshape(-1, 2)
    assert parent[0] == -1
    seen = np.zeros(2 * n - 1, np.int32)
    for i in range(n - 1):
        for c in child[i]:
            assert parent[c] == i
            seen[c] += 1
    assert seen[0] == 0 and (seen[1:] == 1).all()
    # refit boxes contain children
    box = out["bin_box"]
    for i in range(n - 1):
        for c in child[i]:
            assert (box[i, :3] <= box[c, :3]).all() and (box[i, 3:] >= box[c, 3:]).all()


def test_lbvh_restatement_invariants(built):
    P, idx = S.heightfield_mesh(12, 9, None, seed=2)
    ps = O.PortScene()
    ps.add_mesh(P, idx)
    box = ps.tri_boxes()
    n = len(box)
    out = O.lbvh_build(box)
    assert (np.diff(out["morton"].astype(np.int64)) >= 0).all()
    assert sorted(out["prim_order"].tolist()) == list(range(n))
    _check_tree(out, n)
    # every prim slot is referenced by exactly one wide leaf
    wide = out["wide_nodes"]
    cover = np.zeros(n, np.int32)
    for w in wide:
        for k in range(4):
            ref = int(w[24 + k])
            if ref == 0x7FFFFFFF:
                continue
            if ref & 0x80000000:
                first, cnt = (ref & 0x7FFFFFFF) >> 3, (ref & 7) + 1
                assert cnt <= 4
                cover[first:first + cnt] += 1
            else:
                assert 0 < ref < len(wide)
    assert (cover == 1).all()


def test_lbvh_restatement_degenerate_inputs(built):
    # all-identical boxes (every Morton code equal -> index tie-break), n = 1, n = 2
    for n in (1, 2, 3, 50):
        box = np.tile(np.array([[0, 0, 0, 1, 1, 1]], np.float32), (n, 1))
        out = O.lbvh_build(box)
        assert len(out["wide_nodes"]) >= 1
        if n > 1:
            _check_tree(out, n)


def test_abi_exports_every_declared_symbol(built):
    """libb2rt.so loads and exports exactly the functions include/b2rt.h declares."""
    from cuda_raytrace_b200 import b2rt
    hdr = open(os.path.join(ROOT, "include", "b2rt.h")).read()
    declared = set(re.findall(r"\b(b2rt_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(b2rt.SYMBOLS)
    lib = ctypes.CDLL(b2rt.LIB_PATH)
    for s in declared:
        assert hasattr(lib, s), s


def test_no_cpu_fallback(built):
    """Without a GPU the product path must fail loudly, never route to the oracle."""
    from cuda_raytrace_b200 import b2rt
    lib = b2rt.load()
    if lib.b2rt_device_count() > 0:
        pytest.skip("a GPU is visible")
    with pytest.raises(b2rt.B2rtError) as e:
        b2rt.Context(0)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)
    # and the product package never imports the oracle
    pkg = os.path.join(ROOT, "cuda_raytrace_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                for needle in ("import oracle", "from oracle", "liboracle", "libpbrt_oracle", "oracle_port.h"):
                    assert needle not in src, (f, needle)


def test_bench_parity_block_against_reference():
    """bench.py's self-certification: the step's hits against pbrt BVHAccel, disagreements resolved by the exhaustive
    loop.  Here the 'GPU' side is the C port's exhaustive loop (the GPU parity tests prove GPU == port)."""
    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(os.path.dirname(__file__), "..", "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    from cuda_raytrace_b200 import scenes as S
    P, idx = S.heightfield_mesh(40, 40, None, seed=5)
    rays = np.concatenate([S.pinhole_rays(64, 48), S.incoherent_rays(2000, [-1, -0.3, -1], [1, 0.6, 1])])
    port = O.PortScene()
    port.add_mesh(P, idx)
    t, prim, b1, b2 = port.intersect(rays, nthreads=4)
    shadow = S.shadow_rays(rays, t, prim, (0.5, 3.0, 0.5))
    _, sprim, _, _ = port.intersect(shadow, nthreads=4)
    hits = np.zeros((len(rays), 4), np.float32)
    hits[:, 0] = t
    hits[:, 1] = prim.astype(np.int32).view(np.float32)
    cb = bench.cpu_reference_step(P, idx, rays, shadow, 2, sample=None)
    ref = cb.pop("_hits")
    par = bench.hit_parity(hits, (sprim >= 0).astype(np.uint8), ref)
    ref["close"]()
    assert par["ok"] and par["occluded_mismatch"] == 0 and par["primary_rays"] == len(rays), par
    # corrupt one hit distance: the exhaustive loop must side against the 'GPU'
    hits[np.nonzero(prim >= 0)[0][0], 0] *= 2
    cb = bench.cpu_reference_step(P, idx, rays, shadow, 2, sample=None)
    ref = cb.pop("_hits")
    par = bench.hit_parity(hits, (sprim >= 0).astype(np.uint8), ref)
    ref["close"]()
    assert not par["ok"] and par["t_mismatch"] >= 1, par


def test_partial_spheres_match_pbrt_golden(built):
    """zmin / zmax / phimax clipping of Sphere::Intersect and Disk::Intersect (shapes/sphere.cpp:84-106, disk.cpp:73-76): the
    C port against pbrt's own hits (golden from tests/golden/make_golden_partial.py), t bit for bit; the clipped-away part lets
    rays through to what is behind.  The azimuth test is the side-of-the-cut form (no atan2f): a ray within rounding of a
    cut may differ from pbrt."""
    import make_golden_partial as GP
    g = np.load(os.path.join(ROOT, "tests", "golden", "partial_spheres.npz"))
    ps = O.PortScene()
    GP.fill(ps, "port")
    r = GP.rays()
    t, prim, _, _ = ps.intersect(r, 4)
    same = t.view(np.uint32) == g["t"].view(np.uint32)
    assert (~same).sum() <= 4, int((~same).sum())        # BVHAccel's unpadded boxes may drop a hit; a ray on a phimax cut
    assert (prim[same] != g["prim"][same]).sum() <= 2
    assert np.array_equal(ps.occluded(r, 4)[same], g["occ"][same])
    full = O.PortScene()
    (Pg, ig), sph = GP.shapes()
    full.add_mesh(Pg, ig)
    for (o2w, w2o, rad, z0, z1, pm) in sph:
        full.add_sphere(o2w, w2o, rad)
    full.add_disk(*GP.disk()[:5])
    tf, pf, _, _ = full.intersect(r, 4)
    assert (tf != t).sum() > 300                                          # the clipping changes many hits
