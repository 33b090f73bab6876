"""GPU parity tests (-m gpu) for final gathering (pbrt photonmap with finalgather true; csrc/finalgather.cu), every call
through the C ABI.  Bars: direct / radiance photons, irradiance (EPhoton), Lo of the radiance photons and WHICH radiance
photon a lookup returns bit-exact against the C oracle (oracle/oracle_fg.h, itself pinned to pbrt's own EPhoton and
KdTree<RadiancePhoton> lookup); the gather's radiance within 1e-5 relative of the oracle; a rendered image against the
reference's own CPU renderer with finalgather true (tests/golden/make_golden_images_fg.py) within Monte-Carlo bars."""
import os
import numpy as np
import pytest

from helpers import GOLDEN_DIR, load_golden
from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S
import make_golden_photon as G
import make_golden_fg as F

pytestmark = pytest.mark.gpu
NUSED, MAXD2 = 50, np.float32(0.3) * np.float32(0.3)
INV_PI = np.float32(0.31830988618379067154)
# cos(Radians(10)) as pbrt forms it (photonmap.cpp:296, core/pbrt.h Radians): ((float)M_PI / 180.f) * deg, cos in double
COS10 = np.float32(np.cos(np.float64((np.float32(np.pi) / np.float32(180.0)) * np.float32(10.0))))


def _raw(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def fctx(built):
    c = b2rt.Context(0, seed=G.SEED)
    S.fill_scene(c, S.cornell_scene())
    c.build_accel()
    c.photons_trace(0, F.FG_PATHS)
    c.photons_trace_fg(0, F.FG_PATHS)
    for m in (0, 1, 2):
        c.grid_build(m)
    yield c
    c.close()


@pytest.fixture(scope="module")
def oracle_maps():
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    c, i = sc.photon_trace(G.SEED, 0, F.FG_PATHS)
    d, r = sc.photon_trace_fg(G.SEED, 0, F.FG_PATHS)
    return sc, {0: c, 1: i, 2: d, 3: r}


def test_fg_photons_bit_exact(fctx, oracle_maps):
    _, om = oracle_maps
    for m in (2, 3):
        n, paths = fctx.photons_count(m)
        assert (n, paths) == (len(om[m]), F.FG_PATHS)
        assert np.array_equal(_raw(fctx.photons_download(m)), _raw(om[m]))
    # the paths are the photon pass's own: maps 0 and 1 are untouched by the second pass
    assert np.array_equal(_raw(fctx.photons_download(1)), _raw(om[1]))


def test_fg_deposit_limits_follow_pbrt(built):
    """pbrt's depositedPhoton (photonmap.cpp:455-490): with limits (caustic_paths, indirect_paths) a caustic-type hit makes a
    radiance photon only on paths below the first, direct / indirect hits (and the direct photon) only below the second."""
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    n = 30000
    with b2rt.Context(0, seed=G.SEED) as c:
        S.fill_scene(c, S.cornell_scene())
        c.build_accel()
        for lim in ((12000, 20000), (30000, 9000), (0, 0)):
            c.photons_clear()
            c.fg_deposit_limits(*lim)
            c.photons_trace_fg(0, n)
            d, r = sc.photon_trace_fg(G.SEED, 0, n, limits=lim)
            assert np.array_equal(_raw(c.photons_download(2)), _raw(d)) and np.array_equal(_raw(c.photons_download(3)), _raw(r))
            assert (d["path"] < lim[1]).all() and (r["path"] < max(lim)).all()
            assert c.photons_count(2)[1] == min(n, lim[1])            # nDirectPaths: paths shot while the indirect map collected
        c.fg_deposit_limits(-1, -1)
        c.photons_clear()
        c.photons_trace_fg(0, n)
        d0, r0 = sc.photon_trace_fg(G.SEED, 0, n)
        assert np.array_equal(_raw(c.photons_download(3)), _raw(r0)) and len(r0) > len(r)


def test_fg_photons_independent_of_path_range_split(built, oracle_maps):
    _, om = oracle_maps
    with b2rt.Context(0, seed=G.SEED) as c:
        S.fill_scene(c, S.cornell_scene())
        c.build_accel()
        for a, b in ((0, 3001), (3001, 12000), (12000, F.FG_PATHS)):
            c.photons_trace_fg(a, b - a)
        for m in (2, 3):
            assert np.array_equal(_raw(c.photons_download(m)), _raw(om[m]))


@pytest.mark.parametrize("m", [1, 2])
def test_irradiance_bit_exact(fctx, oracle_maps, m):
    _, om = oracle_maps
    g = O.grid_build(om[m])
    q = G.queries(300, seed=21)
    for k, md2 in ((50, 0.09), (8, 1e-3), (100, 4.0)):
        E = fctx.irradiance(m, q.view(b2rt.QUERY_DTYPE).reshape(-1), k, md2)
        ref = O.irradiance(g["sorted"], F.FG_PATHS, q, k, md2)
        assert np.abs(ref).max() > 0
        assert np.array_equal(_raw(E), _raw(ref))


def test_irradiance_matches_pbrt_golden(built):
    """EPhoton of the reference itself (golden) on the 100 000-path indirect map."""
    with b2rt.Context(0, seed=G.SEED) as c:
        S.fill_scene(c, S.cornell_scene())
        c.build_accel()
        c.photons_trace(0, G.NPATHS)
        c.grid_build(1)
        for case, (k, md2) in enumerate(F.E_CASES):
            E = c.irradiance(1, G.queries().view(b2rt.QUERY_DTYPE).reshape(-1), k, md2)
            ref = load_golden("fg")["ephoton_%d" % case]
            assert np.abs(E - ref).max() <= 2e-6 * np.abs(ref).max()


@pytest.fixture(scope="module")
def computed(fctx, oracle_maps):
    """Lo of every radiance photon: GPU (b2rt_radiance_compute) and oracle (EPhoton sums in pbrt's order)."""
    _, om = oracle_maps
    rad = om[3]
    q = np.zeros((len(rad), 12), np.float32)
    q[:, 0:3], q[:, 3:6] = rad["p"], O.oct_decode(rad["wi_oct"])
    E = np.zeros((len(rad), 3), np.float32)
    for m in (2, 1, 0):                                       # direct + indirect + caustic (photonmap.cpp:600-605)
        if len(om[m]):
            E = E + O.irradiance(O.grid_build(om[m])["sorted"], F.FG_PATHS, q, NUSED, float(MAXD2))
    lo = rad.copy()
    lo["alpha"] = (INV_PI * rad["alpha"]) * E
    fctx.radiance_compute(NUSED, 0.3)
    return lo


def test_radiance_compute_bit_exact(fctx, computed):
    got = fctx.photons_download(3)
    assert np.abs(computed["alpha"]).max() > 0
    assert np.array_equal(_raw(got), _raw(computed))


def test_radiance_lookup_exact(fctx, computed):
    g = fctx.grid_download(3)
    og = O.grid_build(computed)
    assert np.array_equal(_raw(g["sorted"]), _raw(og["sorted"]))
    q = np.concatenate([F.lookup_queries(600, seed=11), F.lookup_queries(400, seed=13)])
    rgb, idx = fctx.radiance_lookup(q.view(b2rt.QUERY_DTYPE).reshape(-1))
    orgb, oidx = O.radiance_lookup(og["sorted"], q)
    assert np.array_equal(idx, oidx) and np.array_equal(_raw(rgb), _raw(orgb))
    # and it is the photon the reference's own kd-tree lookup returns (golden, indices in path order)
    gold = load_golden("fg")["lookup_index"]
    mine = np.where(idx[:600] >= 0, g["order"][np.maximum(idx[:600], 0)].astype(np.int64), -1)
    assert np.array_equal(mine, gold)


def _gather_queries(n, seed):
    q = G.queries(n, seed=seed)
    rng = np.random.default_rng(seed)
    q[:, 9] = 1e-3
    q[:, 10] = rng.integers(0, 1 << 20, n).astype(np.uint32).view(np.float32)
    q[:, 11] = (rng.integers(0, 64, n).astype(np.uint32) | np.uint32(1 << 22)).view(np.float32)
    return q


@pytest.mark.parametrize("gs", [32, 2, 100])
def test_final_gather_matches_oracle(fctx, oracle_maps, computed, gs):
    sc, om = oracle_maps
    ind = O.grid_build(om[1])["sorted"]
    rad = O.grid_build(computed)["sorted"]
    q = _gather_queries(96, seed=31)
    fctx.set_final_gather(True, gs, 10.0)
    got = fctx.final_gather(q.view(b2rt.QUERY_DTYPE).reshape(-1), float(MAXD2))
    cos_angle = COS10
    ref = sc.final_gather(ind, rad, q, gs, cos_angle, G.SEED)
    fctx.set_final_gather(False)
    assert np.abs(ref).max() > 0 and np.isfinite(got).all()
    err = np.abs(got - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-30)
    assert err.max() <= 1e-5, (err.max(), float((_raw(got) == _raw(ref)).mean()))


@pytest.mark.parametrize("variant", os.environ.get("B2RT_TEST_FG_VARIANTS", "0,3").split(","))
def test_final_gather_lookup_variants_are_bit_identical(fctx, computed, monkeypatch, variant):
    """B2RT_FG_VARIANT: 3 (default) traces the gather rays, sorts their hits by radiance-grid cell, looks the radiance
    photons up over the sorted hits and combines per shading point in three launches; 0 is the fused warp-per-shading-point
    kernel.  Every ray's arithmetic and the order of the final sum are the same, so the output must not change by a bit,
    whichever variant is the default (also with gather-sample counts that are not one warp)."""
    q = np.concatenate([_gather_queries(96, seed=31), _gather_queries(160, seed=32)])
    # shading points whose gather rays mostly end on the glass / mirror spheres (far lookups): on the floor between them
    q[96:, 0] = np.linspace(-0.8, 0.8, 160, dtype=np.float32)
    q[96:, 1] = 0
    q[96:, 2] = np.linspace(-0.6, 0.6, 160, dtype=np.float32)
    q[96:, 3:6] = (0, 1, 0)
    fctx.set_final_gather(True, 32, 10.0)
    ref = fctx.final_gather(q.view(b2rt.QUERY_DTYPE).reshape(-1), float(MAXD2))
    fctx.set_final_gather(False)
    monkeypatch.setenv("B2RT_FG_VARIANT", variant)
    with b2rt.Context(0, seed=G.SEED) as c:
        S.fill_scene(c, S.cornell_scene())
        c.build_accel()
        c.photons_trace(0, F.FG_PATHS)
        c.photons_trace_fg(0, F.FG_PATHS)
        for m in (0, 1, 2):
            c.grid_build(m)
        c.radiance_compute(NUSED, 0.3)
        c.set_final_gather(True, 32, 10.0)
        got = c.final_gather(q.view(b2rt.QUERY_DTYPE).reshape(-1), float(MAXD2))
        assert np.array_equal(_raw(c.photons_download(3)), _raw(computed))
    assert np.abs(ref).max() > 0 and np.array_equal(_raw(got), _raw(ref))


def test_final_gather_small_map_radius_doubling(built):
    """fewer than 50 indirect photons and queries far from them: the doubling search ends with every photon's direction"""
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    with b2rt.Context(0, seed=7) as c:
        S.fill_scene(c, S.cornell_scene())
        c.build_accel()
        c.photons_trace(0, 40)
        c.photons_trace_fg(0, 40)
        n_ind = c.photons_count(1)[0]
        assert 0 < n_ind < 50
        for m in (0, 1, 2):
            if c.photons_count(m)[0]:
                c.grid_build(m)
        c.radiance_compute(NUSED, 0.3)
        c.set_final_gather(True, 32, 10.0)
        q = _gather_queries(40, seed=5)
        got = c.final_gather(q.view(b2rt.QUERY_DTYPE).reshape(-1), 1e-4)
        ind = c.grid_download(1)["sorted"]
        rad = c.grid_download(3)["sorted"]
    cos_angle = COS10
    ref = sc.final_gather(ind, rad, q, 32, cos_angle, 7)
    err = np.abs(got - ref).max(axis=1) / np.maximum(np.abs(ref).max(axis=1), 1e-30)
    assert np.abs(ref).max() > 0 and err.max() <= 1e-5, err.max()


def _block(img, b):
    h, w, _ = img.shape
    return img.reshape(h // b, b, w // b, b, 3).mean((1, 3))


def _fg_frame(c, cam, prm, maxd, spp, **kw):
    c.photons_shoot(int(prm["ncaustic"]), int(prm["nindirect"]), batch_paths=65536)
    c.photons_trace_fg(0, c.photons_count(1)[1])
    for m in (0, 1, 2):
        c.grid_build(m)
    c.radiance_compute(int(prm["nused"]), maxd)
    c.set_final_gather(True, int(prm["gs"]), kw.get("angle", 10.0))
    return c.render_frame(cam, spp=spp, nused=int(prm["nused"]), max_dist=maxd)


def test_fg_image_matches_pbrt_golden(built):
    g = np.load(os.path.join(GOLDEN_DIR, "images_cornell_fg.npz"))
    ref = g["fg"]
    res, prm = ref.shape[0], {k: g["params"][n] for n, k in enumerate(("spp", "ncaustic", "nindirect", "nused", "gs"))}
    with b2rt.Context(0, seed=11) as c:
        S.fill_scene(c, S.cornell_scene(boxes=True))
        c.build_accel()
        cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, res, res)
        img = _fg_frame(c, cam, prm, float(g["maxdist"]), 256, angle=float(g["gatherangle"]))
        # tiling independence with final gathering on: a band of the frame is bit-identical to the whole frame's rows
        band = c.render_frame(cam, spp=4, nused=int(prm["nused"]), max_dist=float(g["maxdist"]), window=(0, res, 16, 32))
        whole = c.render_frame(cam, spp=4, nused=int(prm["nused"]), max_dist=float(g["maxdist"]))
    assert np.array_equal(_raw(band), _raw(whole[16:32]))
    m = S.image_metrics(_block(img, 8), _block(ref, 8))
    # both sides are Monte-Carlo estimates over different photon sets (pbrt: Halton + MT19937, ours: Philox); the bars are
    # loose enough for that and tight enough to catch a wrong weight in the gather (the indirect term is ~1/3 of the image)
    assert abs(m["bias_rel"]) < 0.03, m
    assert m["mean_rel"] < 0.06, m
