"""Every kernel schedule must return the same bytes (-m gpu).  The schedule is chosen when a context is created
(environment: B2RT_TRACE_VARIANT / B2RT_TRACE_REFILL, B2RT_PHOTON_VARIANT, B2RT_GATHER_VARIANT), so each case makes its
own context.  Traversal is also checked against the CPU oracle; photon and gather variants against each other and,
through test_gpu_photon.py (default variants), against the oracle."""
import os

import numpy as np
import pytest

from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S

pytestmark = pytest.mark.gpu
_KEYS = ("B2RT_TRACE_VARIANT", "B2RT_TRACE_REFILL", "B2RT_TRACE_CHUNK", "B2RT_PHOTON_VARIANT", "B2RT_GATHER_VARIANT",
         "B2RT_TRACE_PLANES", "B2RT_TRACE_SORT")


def _ctx(seed, **env):
    old = {k: os.environ.get(k) for k in _KEYS}
    try:
        for k in _KEYS:
            os.environ.pop(k, None)
        for k, v in env.items():
            os.environ[k] = str(v)
        return b2rt.Context(0, seed=seed)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _raw(a):
    return np.ascontiguousarray(a).view(np.uint8)


@pytest.fixture(scope="module")
def terrain(built):
    P, idx = S.heightfield_mesh(96, 96, None, seed=11)
    lo, hi = P.min(0) - 0.05, P.max(0) + 0.3
    rays = np.concatenate([S.pinhole_rays(160, 90), S.incoherent_rays(30000, lo, hi, seed=5)])
    port = O.PortScene()
    port.add_mesh(P, idx)
    t, prim, b1, b2 = port.intersect(rays, nthreads=8)
    return P, idx, rays, (t, prim, b1, b2)


@pytest.mark.parametrize("env", [
    dict(B2RT_TRACE_VARIANT=0),
    dict(B2RT_TRACE_VARIANT=1), dict(B2RT_TRACE_VARIANT=1, B2RT_TRACE_REFILL=20),
    dict(B2RT_TRACE_VARIANT=2), dict(B2RT_TRACE_VARIANT=2, B2RT_TRACE_REFILL=12),
    dict(B2RT_TRACE_VARIANT=2, B2RT_TRACE_REFILL=32, B2RT_TRACE_CHUNK=96),
    dict(B2RT_TRACE_VARIANT=2, B2RT_TRACE_PLANES=0), dict(B2RT_TRACE_VARIANT=2, B2RT_TRACE_PLANES=2, B2RT_TRACE_SORT=1),
    dict(B2RT_TRACE_VARIANT=2, B2RT_TRACE_PLANES=1, B2RT_TRACE_SORT=0),
], ids=lambda e: "-".join("%s%s" % (k.split("_")[-1].lower(), v) for k, v in e.items()))
def test_traversal_variants_match_oracle(terrain, env):
    P, idx, rays, (t, prim, b1, b2) = terrain
    with _ctx(1, **env) as ctx:
        ctx.add_mesh(P, idx)
        ctx.build_accel()
        hits = ctx.trace_closest(rays)
        occ = ctx.trace_any(rays)
    assert np.array_equal(hits["prim"], prim)
    assert np.array_equal(hits["t"].view(np.uint32), t.view(np.uint32))
    hit = prim >= 0
    assert np.array_equal(hits["b1"].view(np.uint32)[hit], b1.view(np.uint32)[hit])
    assert np.array_equal(hits["b2"].view(np.uint32)[hit], b2.view(np.uint32)[hit])
    assert np.array_equal(occ, hit.astype(np.uint8))


def _photon_frame(env):
    desc = S.cornell_scene(spheres=True)
    with _ctx(7, **env) as ctx:
        S.fill_scene(ctx, desc)
        ctx.build_accel()
        ctx.photons_trace(0, 150000)
        maps = [ctx.photons_download(w).copy() for w in (0, 1)]
        ctx.grid_build(0); ctx.grid_build(1)
        rng = np.random.default_rng(3)
        pts = maps[1][rng.integers(0, len(maps[1]), 3000)]
        q = np.zeros((3000, 12), np.float32)
        q[:, 0:3] = np.stack([pts["p"][:, 0], pts["p"][:, 1], pts["p"][:, 2]], 1) + rng.normal(0, 0.01, (3000, 3)).astype(np.float32)
        q[:, 3:6] = (0, 1, 0); q[:, 6:9] = 0.5
        qq = q.view(b2rt.QUERY_DTYPE).reshape(-1)
        out = []
        for which, k, r2 in ((1, 50, 0.09), (1, 8, 1e-4), (0, 64, 0.01), (1, 0, 0.0009)):
            rgb, rad2, found = ctx.gather(which, qq, k, r2)
            out.append((rgb.copy(), rad2.copy(), found.copy()))
        sets = ctx.gather_sets(1, qq, 50, 0.09).copy()
    return maps, out, sets


def test_photon_and_gather_variants_identical(built):
    base_maps, base_out, base_sets = _photon_frame(dict(B2RT_PHOTON_VARIANT=0, B2RT_GATHER_VARIANT=0))
    for gv in (1, 2):
        _check_same(base_maps, base_out, base_sets, *_photon_frame(dict(B2RT_PHOTON_VARIANT=1, B2RT_GATHER_VARIANT=gv)))
    assert (base_out[0][2] == 50).mean() > 0.5 and (base_out[1][2] < 8).any()    # both full and starved queries occur


def _check_same(base_maps, base_out, base_sets, maps, out, sets):
    for a, b in zip(base_maps, maps):
        assert len(a) > 1000 and np.array_equal(_raw(a), _raw(b))
    assert np.array_equal(base_sets, sets)
    for n, ((rgb0, r0, f0), (rgb1, r1, f1)) in enumerate(zip(base_out, out)):
        assert np.array_equal(f0, f1) and np.array_equal(r0.view(np.uint32), r1.view(np.uint32))
        if n < 3:       # k-nearest: same set, same summation order -> same bits
            assert np.array_equal(rgb0.view(np.uint32), rgb1.view(np.uint32))
        else:           # fixed radius: per-lane accumulation order differs between the schedules
            assert np.allclose(rgb0, rgb1, rtol=1e-5, atol=1e-9)
