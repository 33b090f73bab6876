"""GPU parity tests (-m gpu): every call goes through the C ABI (libb2rt.so via ctypes).
Bars: bit-exact for hit primitive ids, hit distances (the primitive tests reproduce pbrt's arithmetic),
BVH topology, Morton order; occlusion flags exact."""
import numpy as np
import pytest
import torch

from helpers import CASES, load_golden, case_scene, fill, fill_ctx, bits
from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S

pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


# ---- K2: radix sort -------------------------------------------------------------------------------
@pytest.mark.parametrize("n,bits_", [(1, 32), (2, 32), (31, 8), (4096, 16), (4097, 32), (100_003, 30),
                                     (1_000_000, 32), (5_000_000, 24)])
def test_sort_pairs_matches_stable_sort(ctx, n, bits_):
    rng = np.random.default_rng(n)
    eff = 8 * ((bits_ + 7) // 8)
    keys = rng.integers(0, 2 ** eff, n, dtype=np.uint64).astype(np.uint32)
    if n > 1000:
        keys[: n // 3] = keys[n // 2: n // 2 + n // 3]      # plenty of duplicates: stability matters
    vals = np.arange(n, dtype=np.uint32)
    dk, dv = _dev(keys.view(np.int32)), _dev(vals.view(np.int32))
    ctx.sort_pairs_device(n, dk.data_ptr(), dv.data_ptr(), bits_)
    ctx.sync()
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(dv.cpu().numpy().view(np.uint32), order.astype(np.uint32))
    assert np.array_equal(dk.cpu().numpy().view(np.uint32), keys[order])
    ek, ev = O.sort_pairs(keys, vals, bits_)
    assert np.array_equal(ev, dv.cpu().numpy().view(np.uint32))


@pytest.mark.parametrize("variant", ["0", "2"])
def test_sort_skewed_keys_both_schedules(built, monkeypatch, variant):
    """B2RT_SORT_VARIANT: 1 = one-sweep (one histogram pass, one look-back scatter launch per digit), 0 = the three-launch
    pass.  Key sets that stress the look-back and the histogram: all keys equal (every tile waits on one digit chain),
    four distinct top digits (a short cell key), already sorted, reversed, and a ragged last tile."""
    monkeypatch.setenv("B2RT_SORT_VARIANT", variant[0])          # 2 = one-sweep for every size
    rng = np.random.default_rng(7)
    with b2rt.Context(0, seed=1) as c:
        for n, kind, bits_ in [(300_001, "equal", 32), (1_000_003, "top4", 26), (262_144, "sorted", 32),
                               (262_145, "reversed", 32), (4_000_000, "cells", 22), (70_000, "byte", 8)]:
            if kind == "equal":
                keys = np.full(n, 0xdeadbeef, np.uint32)
            elif kind == "top4":
                keys = rng.integers(0, 1 << 26, n, dtype=np.uint64).astype(np.uint32)
            elif kind == "sorted":
                keys = np.sort(rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32))
            elif kind == "reversed":
                keys = np.sort(rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32))[::-1].copy()
            elif kind == "cells":
                keys = (rng.normal(0.5, 0.05, n).clip(0, 1) * ((1 << 22) - 1)).astype(np.uint32)       # clustered cell ids
            else:
                keys = rng.integers(0, 256, n, dtype=np.uint64).astype(np.uint32)
            vals = np.arange(n, dtype=np.uint32)
            dk, dv = _dev(keys.view(np.int32)), _dev(vals.view(np.int32))
            c.sort_pairs_device(n, dk.data_ptr(), dv.data_ptr(), bits_)
            c.sync()
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(dv.cpu().numpy().view(np.uint32), order.astype(np.uint32)), (kind, variant)
            assert np.array_equal(dk.cpu().numpy().view(np.uint32), keys[order]), (kind, variant)


# ---- K1/K3/K4: builder topology, bit-exact against the CPU restatement ----------------------------------
def _tri_only_ctx(ctx, P, idx):
    ctx.clear_scene()
    ctx.add_mesh(P, idx)
    return ctx.build_accel()


@pytest.mark.parametrize("shape", [(1, 1), (3, 2), (12, 9), (40, 40), (224, 224)])
def test_lbvh_bit_exact_vs_restatement(ctx, shape):
    P, idx = S.heightfield_mesh(shape[0], shape[1], None, seed=11)
    info = _tri_only_ctx(ctx, P, idx)
    got = ctx.accel_download()
    ps = O.PortScene()
    ps.add_mesh(P, idx)
    exp = O.lbvh_build(ps.tri_boxes())
    n = len(idx)
    assert info.n_prims == n
    assert np.array_equal(got["morton"], exp["morton"])
    assert np.array_equal(got["prim_order"], exp["prim_order"])
    assert np.array_equal(got["leaf_prims"], exp["prim_order"])
    assert np.array_equal(got["bin_parent"], exp["bin_parent"])
    assert np.array_equal(got["bin_child"], exp["bin_child"])
    assert np.array_equal(bits(got["bin_box"]), bits(exp["bin_box"]))
    assert got["wide_nodes"].shape == exp["wide_nodes"].shape
    assert np.array_equal(got["wide_nodes"], exp["wide_nodes"])


def test_lbvh_degenerate_and_tiny(ctx):
    # one triangle; two identical triangles; 50 identical triangles (all Morton codes equal)
    tri = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)
    for copies in (1, 2, 50):
        P = np.tile(tri, (copies, 1))
        idx = np.arange(3 * copies, dtype=np.int32).reshape(-1, 3)
        info = _tri_only_ctx(ctx, P, idx)
        got = ctx.accel_download()
        ps = O.PortScene()
        ps.add_mesh(P, idx)
        exp = O.lbvh_build(ps.tri_boxes())
        assert np.array_equal(got["wide_nodes"], exp["wide_nodes"])
        rays = S.pack_rays(np.array([[0.2, 0.2, 1.0]]), np.array([[0, 0, -1.0]]), 0.0, np.inf)
        h = ctx.trace_closest(rays)
        assert h["prim"][0] == 0 and h["t"][0] == 1.0      # ties -> min prim id
    # empty scene: builds, everything misses
    ctx.clear_scene()
    ctx.build_accel()
    h = ctx.trace_closest(rays)
    assert h["prim"][0] == -1 and np.isinf(h["t"][0])
    assert ctx.trace_any(rays)[0] == 0
    assert len(ctx.trace_closest(np.zeros((0, 8), np.float32))) == 0


def test_build_is_deterministic_across_runs(ctx):
    P, idx = S.cfg2_mesh(100_000)
    _tri_only_ctx(ctx, P, idx)
    a = ctx.accel_download()
    for _ in range(2):
        _tri_only_ctx(ctx, P, idx)
        b = ctx.accel_download()
        for k in a:
            assert np.array_equal(a[k].view(np.uint32) if a[k].dtype == np.float32 else a[k],
                                  b[k].view(np.uint32) if b[k].dtype == np.float32 else b[k]), k


# ---- K5: traversal vs the reference's golden vectors and the oracle ---------------------------------
@pytest.mark.parametrize("name", CASES)
def test_trace_matches_reference_golden(ctx, name):
    g = load_golden(name)
    fill_ctx(ctx, case_scene(name))
    ctx.build_accel()
    rays = g["rays"]
    hits = ctx.trace_closest(rays)
    # distances: bit-identical to pbrt BVHAccel::Intersect on every ray
    assert np.array_equal(bits(hits["t"]), bits(g["bvh_t"]))
    # ids: identical to pbrt's exhaustive loop (min t, min id) where it was run ...
    sel = g["brute_sel"]
    assert np.array_equal(hits["prim"][sel], g["brute_prim"])
    # ... and to pbrt's BVH except on exact-t ties
    assert int((hits["prim"] != g["bvh_prim"]).sum()) <= int(g["nties"])
    miss = hits["prim"] < 0
    assert np.array_equal(miss, g["bvh_prim"] < 0)
    # barycentrics against pbrt's (u,v) = (b1+b2, b2)
    same = (hits["prim"] == g["bvh_prim"]) & (hits["prim"] >= 0) & (hits["b1"] + hits["b2"] != 0)
    tri = same & (g["bvh_uv"][:, 0] == hits["b1"] + hits["b2"])
    assert tri.sum() >= 0.95 * ((hits["b1"] != 0) | (hits["b2"] != 0)).sum()
    # any-hit == pbrt IntersectP
    occ = ctx.trace_any(rays)
    assert np.array_equal(occ, g["occluded"])


def test_trace_bit_exact_vs_oracle_random_scenes(ctx):
    for seed, (nx, nz) in enumerate([(5, 7), (33, 20), (90, 90)]):
        P, idx = S.heightfield_mesh(nx, nz, None, seed=100 + seed)
        ps = O.PortScene()
        ps.add_mesh(P, idx)
        _tri_only_ctx(ctx, P, idx)
        rays = np.concatenate([S.incoherent_rays(3000, [-1.3, -0.5, -1.3], [1.3, 0.8, 1.3], seed=seed),
                               S.pinhole_rays(40, 30)])
        # finite segments too
        rays[::3, 7] = 0.8
        rays[::5, 3] = 0.05
        t, prim, b1, b2 = ps.intersect(rays, 8)
        h = ctx.trace_closest(rays)
        assert np.array_equal(h["prim"], prim)
        assert np.array_equal(bits(h["t"]), bits(t))
        assert np.array_equal(bits(h["b1"]), bits(b1)) and np.array_equal(bits(h["b2"]), bits(b2))
        assert np.array_equal(ctx.trace_any(rays), ps.occluded(rays, 8))


def test_device_pointer_path_equals_host_path(ctx):
    P, idx = S.cfg2_mesh(100_000)
    _tri_only_ctx(ctx, P, idx)
    rays = S.pinhole_rays(320, 180)
    h = ctx.trace_closest(rays)
    d_r = _dev(rays)
    d_h = torch.zeros((len(rays), 4), dtype=torch.float32, device="cuda")
    d_o = torch.zeros(len(rays), dtype=torch.uint8, device="cuda")
    ctx.trace_closest_device(len(rays), d_r.data_ptr(), d_h.data_ptr())
    ctx.trace_any_device(len(rays), d_r.data_ptr(), d_o.data_ptr())
    ctx.sync()
    got = d_h.cpu().numpy().view(h.dtype).reshape(-1)
    assert np.array_equal(got["prim"], h["prim"]) and np.array_equal(bits(got["t"]), bits(h["t"]))
    assert np.array_equal(d_o.cpu().numpy(), (h["prim"] >= 0).astype(np.uint8))


# ---- full-size properties (BASELINE Cfg2: 1 M triangles, 1920x1080) ------------------------------------
def test_cfg2_full_size_properties(ctx):
    P, idx = S.cfg2_mesh()
    assert len(idx) == 1_000_000
    info = _tri_only_ctx(ctx, P, idx)
    assert info.n_prims == 1_000_000
    acc = ctx.accel_download()
    # Morton order sorted; prim order is a permutation; leaf coverage exact
    assert (np.diff(acc["morton"].astype(np.int64)) >= 0).all()
    assert np.array_equal(np.sort(acc["prim_order"]), np.arange(1_000_000, dtype=np.int32))
    refs = acc["wide_nodes"][:, 24:28].reshape(-1)
    leaf = refs[(refs & 0x80000000) != 0]
    first, cnt = (leaf & 0x7FFFFFFF) >> 3, (leaf & 7) + 1
    cover = np.zeros(1_000_001, np.int64)
    np.add.at(cover, first, 1)
    np.add.at(cover, first + cnt, -1)
    assert (np.cumsum(cover)[:-1] == 1).all()
    # primary rays: each hit point re-derived from (prim, b1, b2) lies on the ray at distance t
    rays = S.pinhole_rays(1920, 1080)
    h = ctx.trace_closest(rays)
    ok = h["prim"] >= 0
    assert ok.mean() > 0.4      # 46 % of the 1080p frame sees the terrain or ground quad
    tri = idx[h["prim"][ok]]
    p1, p2, p3 = P[tri[:, 0]].astype(np.float64), P[tri[:, 1]].astype(np.float64), P[tri[:, 2]].astype(np.float64)
    b1, b2 = h["b1"][ok, None].astype(np.float64), h["b2"][ok, None].astype(np.float64)
    ph = (1 - b1 - b2) * p1 + b1 * p2 + b2 * p3
    pr = rays[ok, 0:3].astype(np.float64) + rays[ok, 4:7].astype(np.float64) * h["t"][ok, None].astype(np.float64)
    assert np.abs(ph - pr).max() < 2e-4
    # idempotence + any-hit consistency: occluded <=> closest hit exists
    h2 = ctx.trace_closest(rays)
    assert np.array_equal(h2.view(np.uint32), h.view(np.uint32))
    assert np.array_equal(ctx.trace_any(rays), ok.astype(np.uint8))
    # shadow rays toward the light: a segment clipped just before its own hit point is never blocked by
    # geometry farther than the segment
    sh = S.shadow_rays(rays, h["t"], h["prim"])
    occ = ctx.trace_any(sh)
    hs = ctx.trace_closest(sh)
    assert np.array_equal(occ, (hs["prim"] >= 0).astype(np.uint8))
    assert (hs["t"][hs["prim"] >= 0] <= sh[hs["prim"] >= 0, 7]).all()
    # a sample of rays against the exhaustive oracle (1 M triangles x 300 rays)
    rng = np.random.default_rng(1)
    sel = rng.choice(len(rays), 300, replace=False)
    ps = O.PortScene()
    ps.add_mesh(P, idx)
    t, prim, _, _ = ps.intersect(rays[sel], 8)
    assert np.array_equal(h["prim"][sel], prim) and np.array_equal(bits(h["t"][sel]), bits(t))


# ---- edge cases of the boundary -------------------------------------------------------------------------------------
def test_empty_batches_and_call_order_errors(built):
    from cuda_raytrace_b200 import b2rt
    with b2rt.Context(0, seed=1) as c:
        rays = S.pinhole_rays(4, 4)
        with pytest.raises(b2rt.B2rtError) as e:        # trace before build: B2RT_E_STATE, like the reference's
            c.trace_closest(rays)                       # launch-before-compile being an error rather than garbage
        assert e.value.code == -3
        c.add_mesh(*S.heightfield_mesh(4, 4, None, seed=1))
        c.build_accel()
        assert len(c.trace_closest(rays[:0])) == 0 and len(c.trace_any(rays[:0])) == 0      # empty batch is not an error
        with pytest.raises(b2rt.B2rtError) as e:
            c._ck(c.lib.b2rt_trace_closest(c.h, 5, None, None, b2rt.HOST))
        assert e.value.code == -1
        with pytest.raises(b2rt.B2rtError) as e:        # gather before any grid was built
            c.gather(1, np.zeros(1, b2rt.QUERY_DTYPE), 8, 0.01)
        assert e.value.code in (-1, -3)
        h = c.trace_closest(rays)                       # the context is still usable after the errors
        assert h.shape == (16,)


def test_axis_parallel_and_degenerate_rays_match_oracle(ctx):
    """Zero direction components (1/d = inf in the slab test), rays lying in the plane of box faces, zero-length and
    inverted segments: everything must still equal the exhaustive oracle loop."""
    P, idx = S.heightfield_mesh(24, 24, None, seed=9)
    ps = O.PortScene()
    ps.add_mesh(P, idx)
    _tri_only_ctx(ctx, P, idx)
    rng = np.random.default_rng(4)
    n = 6000
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = rng.uniform([-1.1, -0.4, -1.1], [1.1, 1.0, 1.1], (n, 3))
    rays[:, 7] = np.inf
    axis = rng.integers(0, 3, n)
    sign = rng.choice([-1.0, 1.0], n)
    rays[np.arange(n), 4 + axis] = sign                               # exactly axis-parallel
    rays[: n // 3, 0:3] = P[rng.integers(0, len(P), n // 3)]          # origins exactly on vertices (box planes)
    rays[: n // 3, 1] += 0.5
    rays[n // 3: n // 2, 5] = -1.0; rays[n // 3: n // 2, 4] = 0; rays[n // 3: n // 2, 6] = 0    # straight down
    two = rng.uniform(-1, 1, (n // 4, 3)).astype(np.float32)
    two[:, rng.integers(0, 3)] = 0.0                                  # one zero component, two free
    rays[-n // 4:, 4:7] = two
    rays[5::17, 7] = rays[5::17, 3]                                   # zero-length segment
    rays[7::19, 3] = 5.0; rays[7::19, 7] = 1.0                        # tmin > tmax
    cls = np.full(n, "axis-parallel, random origin", dtype=object)
    cls[: n // 3] = "axis-parallel, origin above a vertex"
    cls[n // 3: n // 2] = "straight down, random origin"
    cls[-n // 4:] = "one zero component"
    cls[5::17] = "zero-length segment"
    cls[7::19] = "tmin > tmax"
    t, prim, b1, b2 = ps.intersect(rays, 8)
    h = ctx.trace_closest(rays)
    assert (prim >= 0).sum() > 500
    bad = (h["prim"] != prim) | (bits(h["t"]) != bits(t))
    report = {c: (int(bad[cls == c].sum()), int((cls == c).sum()), int(((prim >= 0) & (h["prim"] < 0) & (cls == c)).sum()))
              for c in sorted(set(cls))}
    assert not bad.any(), "mismatches per class (bad, rays, hits the GPU lost): %s" % report
    assert np.array_equal(ctx.trace_any(rays), ps.occluded(rays, 8))


def test_chunked_host_path_with_ragged_tail(ctx):
    """The host-buffer call is a pipeline over 256 K-ray chunks: a batch that is not a multiple of the chunk must give
    the bytes of the single-launch device path, in the caller's order."""
    P, idx = S.cfg2_mesh(100_000)
    _tri_only_ctx(ctx, P, idx)
    base = S.pinhole_rays(640, 360)
    reps = (2 * 262144 + 777) // len(base) + 1
    rays = np.ascontiguousarray(np.tile(base, (reps, 1))[: 2 * 262144 + 777])
    rays[:, 0] += np.linspace(0, 0.05, len(rays), dtype=np.float32)   # every ray distinct
    h = ctx.trace_closest(rays)
    o = ctx.trace_any(rays)
    d_r = _dev(rays)
    d_h = torch.zeros((len(rays), 4), dtype=torch.float32, device="cuda")
    d_o = torch.zeros(len(rays), dtype=torch.uint8, device="cuda")
    ctx.trace_closest_device(len(rays), d_r.data_ptr(), d_h.data_ptr())
    ctx.trace_any_device(len(rays), d_r.data_ptr(), d_o.data_ptr())
    ctx.sync()
    assert np.array_equal(d_h.cpu().numpy().view(np.uint32).reshape(-1), np.ascontiguousarray(h).view(np.uint32))
    assert np.array_equal(d_o.cpu().numpy(), o)


# ---- partial spheres (pbrt "zmin" / "zmax") ---------------------------------------------------------------------------
def test_partial_spheres_bit_exact_vs_oracle_and_pbrt_golden(built):
    """b2rt_add_sphere_ex / b2rt_add_disk_ex: zmin / zmax / phimax clipping (shapes/sphere.cpp:84-106, disk.cpp:73-76; the
    reference's device programs ignore them).  GPU == C port bit for bit; C port == pbrt's own hits (tests/test_oracle_cpu.py + here)."""
    import make_golden_partial as GP
    r = GP.rays()
    ps = O.PortScene()
    GP.fill(ps, "port")
    et, ep, _, _ = ps.intersect(r, 4)
    with b2rt.Context(0, seed=1) as c:
        GP.fill(c, "b2rt")
        c.build_accel()
        hits = c.trace_closest(r)
        occ = c.trace_any(r)
        s = c.surface_at(r)
        with pytest.raises(b2rt.B2rtError):
            c.add_sphere(S.xf_translate(0, 0, 0)[0], 0.5, zmin=0.7, zmax=0.9)      # nothing left after the clamp
    assert np.array_equal(hits["t"].view(np.uint32), et.view(np.uint32)) and np.array_equal(hits["prim"], ep)
    assert np.array_equal(occ, (ep >= 0).astype(np.uint8))
    g = load_golden("partial_spheres")
    same = hits["t"].view(np.uint32) == g["t"].view(np.uint32)
    assert (~same).sum() <= 4 and (hits["prim"][same] != g["prim"][same]).sum() <= 2
    # rays that enter through the cut see the INSIDE of the sphere: the hit is the far root, the normal still points outwards
    nprims = ep.max() + 1
    inside = (ep >= nprims - 6) & (ep < nprims - 1) & (np.einsum("ij,ij->i", s["n"], r[:, 4:7]) > 0)
    assert inside.sum() > 100


# ---- per-stage device times (b2rt_get_timings) ------------------------------------------------------------------------
def test_stage_timings(built):
    """off by default (no events, no extra synchronisation); when enabled every stage entry point is bracketed by CUDA
    events and b2rt_get_timings returns device milliseconds and call counts per stage"""
    desc = S.cornell_scene()
    rays = S.pinhole_rays(64, 48)
    with b2rt.Context(0, seed=2) as c:
        S.fill_scene(c, desc)
        c.build_accel()
        c.trace_closest(rays)
        assert all(v == (0.0, 0) for v in c.timings().values())
        c.timings_enable(True)
        c.build_accel()
        c.trace_closest(rays); c.trace_closest(rays); c.trace_any(rays)
        c.photons_shoot(500, 20000, batch_paths=16384)          # calls the trace stage inside: counted once, as "photons"
        c.grid_build(1)
        q = np.zeros((32, 12), np.float32); q[:, 0] = np.linspace(-0.8, 0.8, 32); q[:, 4] = 1; q[:, 6:9] = 0.5
        c.gather(1, q.view(b2rt.QUERY_DTYPE).reshape(-1), 20, 0.09)
        cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 32, 32)
        c.render_frame(cam, spp=1, nused=20, max_dist=0.3)
        t = c.timings(reset=True)
        assert [t[k][1] for k in ("build", "trace_closest", "trace_any", "photons", "grid", "gather", "render", "film")] == [1, 2, 1, 1, 1, 1, 1, 0]
        assert all(t[k][0] > 0 for k in ("build", "trace_closest", "trace_any", "photons", "grid", "gather", "render"))
        assert t["build"][0] < 50 and t["trace_closest"][0] < 50
        assert all(v == (0.0, 0) for v in c.timings().values())
        c.timings_enable(False)
        c.trace_closest(rays)
        assert c.timings()["trace_closest"] == (0.0, 0)
