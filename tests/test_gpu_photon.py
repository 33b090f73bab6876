"""GPU parity tests (-m gpu) for the photon pass, the hash grid and the radiance estimate.  Every call goes through
the C ABI (libb2rt.so via ctypes).  Bars: photons, grid keys / order / cell ranges and kNN sets bit-exact against the
C oracle (oracle/oracle_photon.c, itself pinned to pbrt-v2); kNN radiance bit-exact (same summation order);
fixed-radius radiance 1e-5 relative (sum order differs); radiance vs the golden pbrt LPhoton values 1e-5."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S
import make_golden_photon as G

pytestmark = pytest.mark.gpu


def _raw(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def pctx(built):
    c = b2rt.Context(0, seed=G.SEED)
    S.fill_scene(c, S.cornell_scene())
    c.build_accel()
    yield c
    c.close()


@pytest.fixture(scope="module")
def oracle_photons():
    return G.oracle_photons()


def test_photon_trace_bit_exact(pctx, oracle_photons):
    pctx.photons_clear()
    nc, ni = pctx.photons_trace(0, G.NPATHS)
    c, i = pctx.photons_download(b2rt.MAP_CAUSTIC), pctx.photons_download(b2rt.MAP_INDIRECT)
    oc, oi = oracle_photons
    assert (nc, ni) == (len(oc), len(oi))
    assert np.array_equal(_raw(c), _raw(oc))
    assert np.array_equal(_raw(i), _raw(oi))


@pytest.mark.parametrize("splits", [2, 3, 8])
def test_photon_array_independent_of_path_range_split(pctx, oracle_photons, splits):
    """the multi-GPU contract: rank r traces [r*n/G, (r+1)*n/G); concatenation in rank order == the G = 1 array"""
    pctx.photons_clear()
    edges = np.linspace(0, 40000, splits + 1).astype(np.int64)
    for a, b in zip(edges[:-1], edges[1:]):
        pctx.photons_trace(int(a), int(b - a))
    i = pctx.photons_download(b2rt.MAP_INDIRECT)
    pctx.photons_clear()
    pctx.photons_trace(0, 40000)
    assert np.array_equal(_raw(pctx.photons_download(b2rt.MAP_INDIRECT)), _raw(i))
    oc, oi = oracle_photons
    assert np.array_equal(_raw(i), _raw(oi[oi["path"] < 40000]))


def test_flags_and_depth(pctx):
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    for kw in [dict(caustic=True, indirect=False), dict(caustic=False, indirect=True)]:
        for depth in (1, 2, 7):
            pctx.photons_clear()
            pctx.photons_trace(100, 5000, max_depth=depth, **kw)
            oc, oi = sc.photon_trace(G.SEED, 100, 5000, max_depth=depth, **kw)
            assert np.array_equal(_raw(pctx.photons_download(0)), _raw(oc))
            assert np.array_equal(_raw(pctx.photons_download(1)), _raw(oi))


def test_point_light_scene_bit_exact(built):
    desc = S.cornell_scene(spheres=True, boxes=True, light="point")
    sc = O.PortFullScene()
    S.fill_scene(sc, desc)
    with b2rt.Context(0, seed=99) as c:
        S.fill_scene(c, desc)
        c.build_accel()
        c.photons_trace(0, 20000)
        oc, oi = sc.photon_trace(99, 0, 20000)
        assert len(oc) > 50 and len(oi) > 5000
        assert np.array_equal(_raw(c.photons_download(0)), _raw(oc))
        assert np.array_equal(_raw(c.photons_download(1)), _raw(oi))


def test_shoot_follows_pbrt_stopping_rule(pctx):
    (pc, pi), (nc, ni) = pctx.photons_shoot(500, 20000, batch_paths=4096)
    assert nc >= 500 and ni >= 20000
    assert pc % 4096 == 0 and pi % 4096 == 0 and pc != pi         # each map stops counting paths once it is full
    gold = load_golden("photon_cornell")
    assert abs(ni / pi / gold["pbrt_indirect_per_path"] - 1) < 0.03
    assert abs(nc / pc / gold["pbrt_caustic_per_path"] - 1) < 0.15


def test_shoot_gives_up_like_pbrt(built):
    """pbrt's unsuccessful() (photonmap.cpp:182-184, 526-537): after 500 000 paths a wanted map with fewer than
    blockSize / 1024 = 4 photons ends the shooting and every map is erased -- no error, the frame renders without maps.
    A diffuse-only scene can never store a caustic photon; a scene with sparse caustics must NOT give up."""
    with b2rt.Context(0, seed=G.SEED) as c:
        S.fill_scene(c, S.cornell_scene(spheres=False, boxes=True))
        c.build_accel()
        (pc, pi), (nc, ni) = c.photons_shoot(1000, 10_000_000, batch_paths=262144)
        assert (nc, ni) == (0, 0) and "giving up" in c.last_error()
        assert c.photons_count(1)[0] == 0
        cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 16, 16)
        img = c.render_frame(cam, spp=1, nused=20, max_dist=0.3)          # still renders (direct lighting only)
        assert np.isfinite(img).all() and img.max() > 0
        # without a caustic request the same scene fills its indirect map
        (pc, pi), (nc, ni) = c.photons_shoot(0, 20000, batch_paths=65536)
        assert nc == 0 and ni >= 20000
    with b2rt.Context(0, seed=G.SEED) as c:
        # a tiny glass ball: a few caustic photons per 100 000 paths -- pbrt keeps shooting (>= 4 stored), so must we;
        # the old rule (photons < paths shot / 1024) stopped here
        desc = S.cornell_scene(spheres=False)
        desc["shapes"].insert(5, ("sphere", *S.xf_translate(0.0, 0.03, 0.0), 0.03, 3))
        desc["lights"] = [("disk", 6, (17.0, 12.0, 4.0), 4)]
        S.fill_scene(c, desc)
        c.build_accel()
        (pc, pi), (nc, ni) = c.photons_shoot(40, 1000, batch_paths=262144)
        assert nc >= 40 and pc > 500000 and nc < pc // 1024


@pytest.mark.parametrize("cell", [0.0, 0.05, 0.3])
def test_grid_bit_exact(pctx, oracle_photons, cell):
    pctx.photons_clear()
    pctx.photons_trace(0, G.NPATHS)
    for which, ph in ((b2rt.MAP_CAUSTIC, oracle_photons[0]), (b2rt.MAP_INDIRECT, oracle_photons[1])):
        gi = pctx.grid_build(which, cell_size=cell)
        got = pctx.grid_download(which)
        exp = O.grid_build(ph, cell_size=cell)
        assert list(gi.dims) == list(exp["dims"]) and gi.cell_size == np.float32(exp["cell_size"])
        assert np.array_equal(np.array(list(gi.origin), np.float32), exp["origin"])
        assert np.array_equal(got["keys"], exp["keys"])
        assert np.array_equal(got["order"], exp["order"])
        assert np.array_equal(got["cell_start"], exp["cell_start"])
        assert np.array_equal(_raw(got["sorted"]), _raw(exp["sorted"]))


@pytest.mark.parametrize("case", range(len(G.GATHER_CASES)))
@pytest.mark.parametrize("cell", [0.0, 0.11])
def test_gather_matches_oracle_and_pbrt_golden(pctx, oracle_photons, case, cell):
    pctx.photons_clear()
    pctx.photons_trace(0, G.NPATHS)
    pctx.grid_build(b2rt.MAP_INDIRECT, cell_size=cell)
    sorted_ph = pctx.grid_download(b2rt.MAP_INDIRECT)["sorted"]
    k, md2 = G.GATHER_CASES[case]
    q = G.queries()
    rgb, r2, found = pctx.gather(b2rt.MAP_INDIRECT, q.view(b2rt.QUERY_DTYPE).reshape(-1), k, md2)
    ergb, er2, efound = O.gather(sorted_ph, G.NPATHS, q, k, md2)
    assert np.array_equal(found, efound)
    assert np.array_equal(_raw(r2), _raw(er2))
    if k > 0:
        idx = pctx.gather_sets(b2rt.MAP_INDIRECT, q.view(b2rt.QUERY_DTYPE).reshape(-1), k, md2)
        _, _, _, eidx = O.gather(sorted_ph, G.NPATHS, q, k, md2, want_sets=True)
        assert np.array_equal(idx, eidx)
        assert np.array_equal(_raw(rgb), _raw(ergb))
    else:
        assert np.abs(rgb - ergb).max() <= 1e-5 * np.abs(ergb).max()
    ref = load_golden("photon_cornell")["lphoton_%d" % case]          # pbrt's own KdTree + LPhoton
    assert np.abs(rgb - ref).max() <= 1e-5 * np.abs(ref).max()


def test_gather_edge_cases(pctx):
    pctx.photons_clear()
    pctx.photons_trace(0, 3000)
    pctx.grid_build(b2rt.MAP_CAUSTIC)
    sp = pctx.grid_download(b2rt.MAP_CAUSTIC)["sorted"]
    q = np.zeros((5, 12), np.float32)
    q[:, 3:6] = (0, 1, 0); q[:, 6:9] = 1
    q[0, 0:3] = (50, 50, 50)            # far outside the grid
    q[1, 0:3] = (0.4, 0, 0.3)           # under the glass sphere: the caustic
    q[2, 0:3] = (-5, 0, 0)
    q[3, 0:3] = sp["p"][0] if len(sp) else 0
    q[4, 0:3] = (0.4, 1e-3, 0.3)
    for k, md2 in [(1, 1e6), (224, 1e6), (7, 1e-12), (0, 1e6)]:
        rgb, r2, found = pctx.gather(b2rt.MAP_CAUSTIC, q.view(b2rt.QUERY_DTYPE).reshape(-1), k, md2)
        ergb, er2, efound = O.gather(sp, 3000, q, k, md2)
        assert np.array_equal(found, efound) and np.array_equal(_raw(r2), _raw(er2))
        assert np.abs(rgb - ergb).max() <= 1e-5 * max(np.abs(ergb).max(), 1e-30)
    with pytest.raises(b2rt.B2rtError):
        pctx.gather(b2rt.MAP_CAUSTIC, q.view(b2rt.QUERY_DTYPE).reshape(-1), 225, 1.0)
    # empty map: grid over zero photons, every estimate is zero
    pctx.photons_clear()
    pctx.grid_build(b2rt.MAP_CAUSTIC)
    rgb, _, found = pctx.gather(b2rt.MAP_CAUSTIC, q.view(b2rt.QUERY_DTYPE).reshape(-1), 10, 1.0)
    assert not rgb.any() and not found.any()
