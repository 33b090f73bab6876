"""GPU parity tests (-m gpu) for the hit attributes (K6): per-vertex shading normals, uv-derived
(dpdu, dpdv), reversed orientation.  b2rt_surface_at (C ABI) against the C restatement bit for bit and against the golden
pbrt-v2 answers (tests/golden/surface.npz)."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S
import make_golden_surface as G
from test_oracle_surface_cpu import compare

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_surface_attributes_match_port_and_pbrt(built, case):
    rays = G.rays()
    ps = O.PortFullScene()
    G.fill(ps, case)
    want = ps.surface_at(rays)
    with b2rt.Context(0, seed=1) as c:
        G.fill(c, case)
        c.build_accel()
        got = c.surface_at(rays)
    assert got.tobytes() == want.tobytes()                 # GPU == C port, every byte incl. misses
    bad, ties = compare(got, load_golden("surface")[case[0]])
    assert ties <= 4 and not any(bad.values()), (bad, ties)


def test_surface_mixed_scene_keeps_plain_meshes_unchanged(built):
    """a scene where only one mesh carries attributes: the other meshes, spheres and disks must give the very bytes they
    give in a scene without any (the fast path of make_surface)"""
    desc = S.cornell_scene()
    rays = np.concatenate([S.pinhole_rays(64, 64, eye=(0, 1, 3.4), look=(0, 1, 0), fov=40),
                           S.incoherent_rays(3000, [-1, 0, -1], [1, 2, 1], seed=8)])
    P, idx, N, uv = S.sphere_mesh(16, 8, 0.25, (0.0, 1.2, 0.0))
    with b2rt.Context(0, seed=1) as c:
        S.fill_scene(c, desc)
        c.build_accel()
        plain = c.surface_at(rays)
    ps = O.PortFullScene()
    S.fill_scene(ps, desc)
    first = ps.add_mesh(P, idx, N=N, uv=uv)
    want = ps.surface_at(rays)
    with b2rt.Context(0, seed=1) as c:
        S.fill_scene(c, desc)
        assert c.add_mesh(P, idx, N=N, uv=uv) == first
        c.build_accel()
        got = c.surface_at(rays)
    assert got.tobytes() == want.tobytes()
    keep = (got["prim"] < first) & (plain["prim"] == got["prim"])
    assert keep.sum() > 1000 and got[keep].tobytes() == plain[keep].tobytes()
    on = got["prim"] >= first
    assert on.sum() > 50 and (np.abs(got["n"][on] - got["ng"][on]).max(axis=1) > 1e-4).any()


def test_smooth_mesh_renders_like_the_analytic_sphere(built):
    """image-level check of the whole path (camera pass, direct lighting, photons): a tessellated matte sphere mesh WITH
    normals is closer to the analytic sphere than the same mesh WITHOUT (faceted)"""
    cam = b2rt.pbrt_camera((0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40, 96, 96)

    def render(shape):
        desc = S.cornell_scene(spheres=False)
        with b2rt.Context(0, seed=11) as c:
            S.fill_scene(c, desc)
            if shape == "analytic":
                c.add_sphere(S.xf_translate(0.0, 0.5, 0.0)[0], 0.5, material=0)
            else:
                P, idx, N, uv = S.sphere_mesh(20, 10, 0.5, (0.0, 0.5, 0.0))
                c.add_mesh(P, idx, material=0, N=N if shape == "smooth" else None, uv=uv if shape == "smooth" else None)
            c.build_accel()
            c.photons_shoot(0, 60000, batch_paths=65536)
            c.grid_build(1)
            return c.render_frame(cam, spp=64, nused=50, max_dist=0.3)
    a, s, f = render("analytic"), render("smooth"), render("faceted")
    win = (slice(40, 80), slice(30, 66))                     # the ball
    da = np.abs(s[win] - a[win]).mean() / a[win].mean()
    df = np.abs(f[win] - a[win]).mean() / a[win].mean()
    assert np.isfinite(s).all() and da < 0.08 and df > da, (da, df)
