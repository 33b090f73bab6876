"""CPU tests (no GPU) of the final-gathering restatements in oracle/oracle_fg.h and of the host-compiled device arithmetic
(tests/hostemu/fg_emu.cu = csrc/b2rt_fg.cuh built for the host), pinned against
  * the committed golden fixture produced by the reference's own pbrt-v2 code (tests/golden/make_golden_fg.py):
    EPhoton values and WHICH radiance photon KdTree<RadiancePhoton>::Lookup + RadiancePhotonProcess returns,
  * the reference itself, live, where oracle/_ref exists (build container)."""
import ctypes as C
import os
import subprocess
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
from cuda_raytrace_b200 import scenes as S
import make_golden_photon as G
import make_golden_fg as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def fg():
    d, r = F.fg_photons()
    return d, r


@pytest.fixture(scope="module")
def emu():
    so = os.path.join(ROOT, "build", "libfg_emu.so")
    src = os.path.join(ROOT, "tests", "hostemu", "fg_emu.cu")
    hdr = os.path.join(ROOT, "cuda_raytrace_b200", "csrc", "b2rt_fg.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.run(["nvcc", "-O2", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off",
                        "-fmad=false", "-Wno-deprecated-gpu-targets", "-I" + os.path.join(ROOT, "include"), src, "-o", so],
                       check=True, capture_output=True)
    lib = C.CDLL(so)
    vp, f = C.c_void_p, C.c_float
    lib.emu_radiance_nearest.argtypes = [vp, f, f, f, vp, vp, vp, vp, C.c_longlong, vp, vp, vp]
    lib.emu_radiance_nearest.restype = None
    lib.emu_radiance_nearest_layout.argtypes = [vp, f, f, f, vp, vp, vp, vp, C.c_longlong, vp, vp, vp, C.c_longlong, C.c_int]
    lib.emu_radiance_nearest_layout.restype = None
    lib.emu_cone_pdf.argtypes = [f]
    lib.emu_cone_pdf.restype = f
    lib.emu_sample_cone.argtypes = [f, f, f, vp, vp]
    lib.emu_sample_cone.restype = None
    lib.emu_oct_encode.argtypes = [vp]
    lib.emu_oct_encode.restype = C.c_uint32
    return lib


def test_fg_paths_are_the_photon_pass_paths(fg):
    """Direct photons are the first diffuse hits of the same paths; radiance photons sit on deposits, ~12.5 % of them."""
    d, r = fg
    sc = O.PortFullScene()
    S.fill_scene(sc, S.cornell_scene())
    c, i = sc.photon_trace(G.SEED, 0, F.FG_PATHS)
    assert np.all(np.diff(d["path"].astype(np.int64)) > 0)                  # at most one direct photon per path, path order
    assert np.all(np.diff(r["path"].astype(np.int64)) >= 0)
    deposits = np.concatenate([d["p"], c["p"], i["p"]])
    dep = {p.tobytes() for p in deposits}
    assert all(p.tobytes() in dep for p in r["p"])                            # every radiance photon lies on a deposit
    n = len(deposits)
    assert abs(len(r) - 0.125 * n) < 4 * np.sqrt(n * 0.125 * 0.875)
    nr = O.oct_decode(r["wi_oct"])
    assert np.allclose(np.linalg.norm(nr, axis=1), 1, atol=1e-6)
    # a split of the path range does not change the photons (multi-GPU sharding of the pass)
    d2a, r2a = sc.photon_trace_fg(G.SEED, 0, 7000)
    d2b, r2b = sc.photon_trace_fg(G.SEED, 7000, F.FG_PATHS - 7000)
    assert np.concatenate([d2a, d2b]).tobytes() == d.tobytes() and np.concatenate([r2a, r2b]).tobytes() == r.tobytes()


@pytest.mark.parametrize("case", range(len(F.E_CASES)))
def test_irradiance_matches_pbrt_ephoton_golden(case):
    _, ind = G.oracle_photons()
    g = O.grid_build(ind)
    k, md2 = F.E_CASES[case]
    E = O.irradiance(g["sorted"], G.NPATHS, G.queries(), k, md2)
    ref = load_golden("fg")["ephoton_%d" % case]
    assert np.abs(ref).max() > 0
    assert np.abs(E - ref).max() <= 2e-6 * np.abs(ref).max()


def test_radiance_lookup_finds_pbrts_photon_golden(fg):
    _, rad = fg
    gold = load_golden("fg")
    assert int(gold["n_radiance"]) == len(rad)
    _, idx = O.radiance_lookup(rad, F.lookup_queries())
    assert (idx >= 0).mean() > 0.9 and np.array_equal(idx, gold["lookup_index"])      # -1: no photon faces the query


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref only exists in the build container")
def test_fg_pieces_match_pbrt_live(fg):
    d, rad = fg
    rs = O.RefPhotonScene()
    q = F.lookup_queries(300, seed=5)
    ref = rs.radiance_lookup(F.radiance9(rad), q)[:, 0].astype(np.int32) - 1
    assert np.array_equal(O.radiance_lookup(rad, q)[1], ref)
    g = O.grid_build(d)
    E = O.irradiance(g["sorted"], F.FG_PATHS, G.queries(200, seed=9), 50, 0.04)
    Er = rs.ephoton(O.photons_to9(g["sorted"]), F.FG_PATHS, G.queries(200, seed=9), 50, 0.04)
    assert np.abs(E - Er).max() <= 2e-6 * np.abs(Er).max()


def _block_sum4(a):
    """sum over 4 x 4 x 4 blocks of a (z, y, x) array (the grid's coarse table, csrc/photon.cu grid_coarse_kernel)"""
    pad = [(0, (-n) % 4) for n in a.shape]
    a = np.pad(a, pad)
    z, y, x = a.shape
    return a.reshape(z // 4, 4, y // 4, 4, x // 4, 4).sum((1, 3, 5))


def _occupancy(cell_start, dims):
    cnt = np.diff(cell_start.astype(np.int64)).reshape(int(dims[2]), int(dims[1]), int(dims[0]))
    c1 = _block_sum4(cnt)
    c2 = _block_sum4(c1)
    return np.ascontiguousarray(c1, np.uint32), np.ascontiguousarray(c2, np.uint32)


def _emu_nearest(emu, g, queries, slack=None, packed=False):
    lo, dims = np.asarray(g["origin"], np.float32), np.asarray(g["dims"], np.int32)
    hi = lo + dims * np.float32(g["cell_size"])
    if slack is None:
        slack = 1e-5 * float(max(np.abs(lo).max(), np.abs(hi).max())) + 1e-4 * g["cell_size"]
    srt = np.ascontiguousarray(g["sorted"])
    cs = np.ascontiguousarray(g["cell_start"], np.uint32)
    q = np.ascontiguousarray(queries, np.float32)
    idx, d2 = np.zeros(len(q), np.int32), np.zeros(len(q), np.float32)
    c1, _ = _occupancy(cs, dims)
    emu.emu_radiance_nearest_layout(lo.ctypes.data, g["cell_size"], g["inv_cell"], slack, dims.ctypes.data, srt.ctypes.data,
                                    cs.ctypes.data, c1.ctypes.data, len(q), q.ctypes.data, idx.ctypes.data, d2.ctypes.data,
                                    len(srt), 1 if packed else 0)
    return idx, d2


@pytest.mark.parametrize("target", [0.0, 1.0, 64.0])
def test_device_shell_search_equals_brute_force(emu, fg, target):
    """csrc/b2rt_fg.cuh fg_radiance_nearest, compiled for the host, over the oracle's grid: the exact nearest facing photon
    for every query -- inside, on the edge of and far outside the grid, for fine and coarse cells."""
    _, rad = fg
    g = O.grid_build(rad, 0.0, target)
    q = np.concatenate([F.lookup_queries(600, seed=11), F.lookup_queries(300, seed=12)])
    q[-20:, 0:3] = rad["p"][:20]                                    # queries exactly on photons
    q[-40:-20, 0:3] = g["origin"] + np.float32(g["cell_size"]) * np.arange(20)[:, None].astype(np.float32)   # on cell faces
    idx, _ = _emu_nearest(emu, g, q)
    _, ref = O.radiance_lookup(g["sorted"], q)
    assert np.array_equal(idx, ref)


def test_packed_layout_returns_the_same_photon(emu, fg):
    """The camera pass searches packed copies (16-byte positions + decoded normals) instead of the sorted records."""
    _, rad = fg
    g = O.grid_build(rad, 0.0, 1.0)
    q = np.ascontiguousarray(F.lookup_queries(400, seed=21), np.float32)
    ref, _ = _emu_nearest(emu, g, q)
    idx, _ = _emu_nearest(emu, g, q, packed=True)
    assert np.array_equal(idx, ref)


def test_device_shell_search_small_and_empty(emu):
    ph = np.zeros(3, O.PHOTON_DTYPE)
    ph["p"] = [(0, 0, 0), (1, 0, 0), (0, 0, 0)]
    up = np.array([0.0, 0.0, 1.0], np.float32)
    ph["wi_oct"] = [emu.emu_oct_encode(up.ctypes.data)] * 3
    g = O.grid_build(ph)
    q = np.zeros((3, 12), np.float32)
    q[:, 3:6] = [(0, 0, 1), (0, 0, -1), (0, 0, 1)]
    q[2, 0:3] = (100, 50, -3)
    idx, _ = _emu_nearest(emu, g, q)
    _, ref = O.radiance_lookup(g["sorted"], q)
    assert np.array_equal(idx, ref) and idx[1] == -1 and idx[0] == 0      # duplicate positions: the lower index wins


def test_device_mis_helpers(emu):
    c = np.float32(np.cos(np.radians(10.0)))
    assert abs(emu.emu_cone_pdf(c) - 1.0 / (2 * np.pi * (1 - float(c)))) < 1e-4
    rng = np.random.default_rng(3)
    for _ in range(200):
        z = rng.normal(size=3)
        z = (z / np.linalg.norm(z)).astype(np.float32)
        out = np.zeros(3, np.float32)
        emu.emu_sample_cone(rng.random(), rng.random(), c, z.ctypes.data, out.ctypes.data)
        assert abs(np.linalg.norm(out) - 1) < 1e-5 and float(out @ z) >= float(c) - 1e-6
