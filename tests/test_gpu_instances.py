"""Object instances without flattening (-m gpu): the two-level BVH behind b2rt_object_begin / b2rt_add_instance through
the C ABI.  Replaces the OptiX Transform nodes of cuda_render/cudarender.cpp:38-103; target behaviour is pbrt's
TransformedPrimitive (core/primitive.cpp:95-120).  Bars: t / b1 / b2 / primitive ids / occlusion bit-exact against the
layered oracle (itself pinned to pbrt's ObjectInstance path, tests/test_oracle_instances_cpu.py) and against pbrt's golden
hits; hit attributes within 2e-5 of pbrt's; an exactly representable scene equals its flattened copy bit for bit."""
import numpy as np
import pytest

from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S
import make_golden_instances as G
from helpers import load_golden

pytestmark = pytest.mark.gpu


def _ctx(desc, seed=1):
    c = b2rt.Context(0, seed=seed)
    ids = S.fill_instanced(c, desc)
    info = c.build_accel()
    return c, ids, info


@pytest.mark.parametrize("variant", ["2", "1", "0"])
def test_instanced_hits_bit_exact_vs_oracle_and_pbrt_golden(built, monkeypatch, variant):
    monkeypatch.setenv("B2RT_TRACE_VARIANT", variant)
    desc, r = G.scene(), G.rays()
    ps = O.PortInstancedScene()
    S.fill_instanced(ps, desc, style="port")
    et, ep, eb1, eb2, ei = ps.intersect(r, 4)
    c, ids, info = _ctx(desc)
    with c:
        assert info.n_prims == ps.n_items
        bases = [c.instance_info(k)[0] for k in ids]
        assert bases == ps.instance_bases()
        hits = c.trace_closest(r)
        occ = c.trace_any(r)
    assert np.array_equal(hits["t"].view(np.uint32), et.view(np.uint32))
    assert np.array_equal(hits["prim"], ep)
    assert np.array_equal(hits["b1"].view(np.uint32), eb1.view(np.uint32)) and np.array_equal(hits["b2"].view(np.uint32), eb2.view(np.uint32))
    assert np.array_equal(occ, (ep >= 0).astype(np.uint8))
    # pbrt's own answer (BVHAccel may drop a hit the exhaustive loop finds; see test_oracle_instances_cpu)
    g = load_golden("instances")
    same = hits["t"].view(np.uint32) == g["t"].view(np.uint32)
    assert (~same).sum() <= 2
    inst = np.full(len(r), -1, np.int32)
    for k, b in enumerate(bases):
        n = c_n = len(ps.objects[ps.instances[k][0]].kind)
        inst[(hits["prim"] >= b) & (hits["prim"] < b + n)] = k
    local = np.where(inst >= 0, hits["prim"] - np.asarray(bases, np.int32)[np.maximum(inst, 0)], hits["prim"])
    bad = same & (hits["prim"] >= 0) & ((inst != g["inst"]) | (local != g["prim"]))
    assert bad.sum() <= 3                                   # exact-t ties
    assert (inst >= 0).sum() > 1000


def test_exact_instances_equal_flattened_build(built):
    """integer translations of grid-quantised meshes: the two-level tree and the flattened tree return the same bits"""
    desc = S.instanced_scene(n_inst=9, seed=4, exact=True)
    r = S.pinhole_rays(160, 120, eye=(0.5, 3.0, 8.0), look=(0.0, 0.0, 0.0), fov=50)
    r[:, :3] = np.float32([0.5, 3.0, 8.0])
    c, ids, _ = _ctx(desc)
    with c:
        h2 = c.trace_closest(r)
        o2 = c.trace_any(r)
    with b2rt.Context(0, seed=1) as f:
        for sh in S.flatten_instanced(desc):
            f.add_mesh(sh[1], sh[2])
        f.build_accel()
        h1 = f.trace_closest(r)
        o1 = f.trace_any(r)
    ni = len(desc["instances"])
    nw = sum(len(sh[2]) for sh in desc["world"])
    assert np.array_equal(h1["t"].view(np.uint32), h2["t"].view(np.uint32))
    assert np.array_equal(h1["b1"], h2["b1"]) and np.array_equal(h1["b2"], h2["b2"]) and np.array_equal(o1, o2)
    assert np.array_equal(np.where(h2["prim"] >= nw, h2["prim"] - ni, h2["prim"]), h1["prim"])
    assert (h2["prim"] >= nw).sum() > 1000


def test_instanced_hit_attributes_match_pbrt(built):
    """b2rt_surface_at: p, geometric and shading normal, BSDF frame of instanced hits (meshes with per-vertex normals under
    rotations, non-uniform scales and a mirroring instance) against pbrt's TransformedPrimitive + GetShadingGeometry"""
    desc, r = G.scene(), G.rays()
    g = load_golden("instances")
    c, ids, _ = _ctx(desc)
    with c:
        s = c.surface_at(r)
        bases = [c.instance_info(k) for k in ids]
    gs = g["surf"]
    hit = (s["prim"] >= 0) & (s["t"].view(np.uint32) == gs["t"].view(np.uint32))
    assert hit.sum() > 5900
    inst = g["inst"]
    # directions: 3e-4 -- flat triangles take their normal from the record's edges (e1, e2) where pbrt sums dp1 / dp2 of the
    # default-uv solve, which differs by rounding on thin triangles, and the instance transform (scales 0.5 .. 1.6) stretches it
    for f, tol in (("p", 2e-5), ("ng", 3e-4), ("n", 3e-4)):
        d = np.abs(s[f][hit] - gs[f][hit]).max()
        assert d < tol * max(1.0, np.abs(gs[f][hit]).max()), (f, d)
    assert np.allclose(s["ray_eps"][hit], gs["ray_eps"][hit], rtol=1e-6)
    # the (sn, tn) frame follows pbrt's dpdu for triangles (analytic shapes use another tangent, as at the top level)
    is_tri = np.zeros(len(r), bool)
    ps = O.PortInstancedScene(); S.fill_instanced(ps, desc, style="port")
    for k, (b, n, o) in enumerate(bases):
        kinds = np.asarray(ps.objects[o].kind)
        m = hit & (inst == k)
        is_tri[m] = kinds[s["prim"][m] - b] >= 0
    m = hit & is_tri & (inst >= 0)
    assert m.sum() > 800
    for f in ("sn", "tn"):
        d = np.abs(s[f][m] - gs[f][m]).max()
        assert d < 3e-4, (f, d)
    # (u, v) are reported for meshes with attributes only: the smooth ball, first mesh of object 0
    nball = len(desc["objects"][0][0][2])
    mu = np.zeros(len(r), bool)
    for k, (b, n, o) in enumerate(bases):
        if o == 0:
            mu |= hit & (inst == k) & (s["prim"] - b < nball)
    assert mu.sum() > 300
    assert np.abs(s["u"][mu] - gs["u"][mu]).max() < 1e-5 and np.abs(s["v"][mu] - gs["v"][mu]).max() < 1e-5
    assert np.abs(s["n"][mu] - s["ng"][mu]).max() > 1e-2             # and they shade smooth: n is the interpolated normal


def test_instanced_frame_and_photons_match_flattened_scene(built):
    """photon pass + camera pass over the two-level scene against the same scene flattened.  Direct lighting: the same
    random streams and hits that differ only by the rounding of instance space versus world space, so the frames agree
    pixel by pixel up to samples that fell on the other side of an edge.  With photon maps: the tangent frames differ (e1
    transformed versus e1 of the transformed triangle is the same direction, but the analytic shapes' tangents are not), so
    photon paths decorrelate after the first bounce and the comparison is statistical."""
    desc = S.instanced_scene(n_inst=8, seed=9)
    cam = b2rt.pbrt_camera((0.5, 3.0, 7.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 50.0, 96, 72)

    def run(fill):
        with b2rt.Context(0, seed=3) as c:
            fill(c)
            c.add_light(0, (60.0, 60.0, 60.0), p=(0.5, 4.5, 1.0))
            c.build_accel()
            nc, ni = c.photons_trace(0, 200000)
            c.grid_build(0); c.grid_build(1)
            direct = c.render_frame(cam, spp=4, caustic=False, indirect=False)
            full = c.render_frame(cam, spp=4, nused=50, max_dist=0.5)
            return nc, ni, direct, full

    def fill_flat(c):
        for sh in S.flatten_instanced(desc):
            S._add_shape(c, sh, "b2rt")

    a, b = run(lambda c: S.fill_instanced(c, desc)), run(fill_flat)
    assert b[2].mean() > 1e-3
    diff = np.abs(a[2] - b[2]).sum(-1) > 1e-4 * (1.0 + b[2].sum(-1))
    assert diff.mean() < 0.03, diff.mean()
    assert abs(a[1] - b[1]) <= 5.0 * np.sqrt(b[1]), (a[:2], b[:2])   # indirect photons stored: independent draws after a bounce
    ind_a, ind_b = (a[3] - a[2]).mean(), (b[3] - b[2]).mean()
    assert ind_b > 1e-4 and abs(ind_a - ind_b) < 0.1 * ind_b, (ind_a, ind_b)


def test_instance_api_errors(built):
    with b2rt.Context(0, seed=1) as c:
        o = c.object_begin()
        with pytest.raises(b2rt.B2rtError):
            c.object_begin()                                 # objects do not nest
        with pytest.raises(b2rt.B2rtError):
            c.add_instance(o, S.xf_translate(0, 0, 0)[0])    # not inside a definition
        P, idx = S.heightfield_mesh(4, 4, None, seed=1)
        with pytest.raises(b2rt.B2rtError):
            c.add_mesh(P, idx, arealight=0)                  # no area lights inside objects
        c.object_end()
        with pytest.raises(b2rt.B2rtError):
            c.add_instance(o, S.xf_translate(0, 0, 0)[0])    # empty object
        with pytest.raises(b2rt.B2rtError):
            c.object_end()
        o2 = c.object_begin(); c.add_mesh(P, idx); c.object_end()
        k = c.add_instance(o2, S.xf_translate(0, 1, 0)[0])   # w2o computed by the library
        c.build_accel()
        assert c.instance_info(k) == (1, len(idx), o2)
        r = S.pinhole_rays(16, 16, eye=(0.0, 4.0, 0.1), look=(0.0, 0.0, 0.0))
        assert (c.trace_closest(r)["prim"] >= 1).any()
        c.clear_scene()
        c.add_mesh(P, idx); c.build_accel()
        assert (c.trace_closest(r)["prim"] >= 0).any()


def test_scene_cache_round_trip(built, tmp_path):
    """b2rt_scene_save / b2rt_scene_load: a scene with attribute meshes, analytic shapes, objects, instances, materials and a
    light goes to a file and comes back into a fresh context; hits, photon maps and frame are the same bytes"""
    desc = S.instanced_scene(n_inst=6, seed=21)
    cam = b2rt.pbrt_camera((0.5, 3.0, 7.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 50.0, 64, 48)
    r = S.incoherent_rays(20000, [-4, -1, -4], [4, 2, 4], seed=2)
    path = str(tmp_path / "scene.b2rt")

    def run(c):
        c.build_accel()
        h = c.trace_closest(r)
        c.photons_trace(0, 50000)
        ph = c.photons_download(1).tobytes()
        c.grid_build(1)
        return h.tobytes(), ph, c.render_frame(cam, spp=2, nused=20, max_dist=0.5).tobytes()

    with b2rt.Context(0, seed=5) as c:
        red = c.add_material(b2rt.MAT_MATTE, (0.7, 0.2, 0.2))
        S.fill_instanced(c, desc)
        c.add_sphere(S.xf_translate(0.3, 0.6, 2.0)[0], 0.4, material=red, zmax=0.2)
        c.add_light(0, (60.0, 60.0, 60.0), p=(0.5, 4.5, 1.0))
        c.scene_save(path)
        want = run(c)
    with b2rt.Context(0, seed=5) as c:
        c.add_mesh(*S.heightfield_mesh(4, 4, None, seed=1))         # replaced by the load
        c.scene_load(path)
        assert c.scene_counts() == {"items": len(desc["world"][0][2]) + 1 + 6 + 1, "lights": 1, "materials": 1, "objects": 2, "instances": 6}
        assert run(c) == want
        with pytest.raises(b2rt.B2rtError):
            c.scene_load(str(tmp_path / "missing.b2rt"))
        open(str(tmp_path / "bad.b2rt"), "wb").write(open(path, "rb").read()[:1000])
        with pytest.raises(b2rt.B2rtError):
            c.scene_load(str(tmp_path / "bad.b2rt"))


def test_instanced_final_gathering_matches_flattened_scene(built):
    """pbrt's default mode (final gathering) over the two-level tree: the radiance-photon pass, the gather rays and their
    hit attributes all go through the instance-aware instantiations; statistical agreement with the flattened scene"""
    desc = S.instanced_scene(n_inst=8, seed=9)
    cam = b2rt.pbrt_camera((0.5, 3.0, 7.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 50.0, 64, 48)

    def run(fill):
        with b2rt.Context(0, seed=3) as c:
            fill(c)
            c.add_light(0, (60.0, 60.0, 60.0), p=(0.5, 4.5, 1.0))
            c.build_accel()
            c.photons_trace(0, 400000)
            c.grid_build(0); c.grid_build(1)
            nd, nr = c.photons_trace_fg(0, 400000)
            c.grid_build(2)
            c.radiance_compute(50, 0.5)
            c.set_final_gather(True, 16, 10.0)
            return nd, nr, c.render_frame(cam, spp=4, nused=50, max_dist=0.5)

    def fill_flat(c):
        for sh in S.flatten_instanced(desc):
            S._add_shape(c, sh, "b2rt")

    a, b = run(lambda c: S.fill_instanced(c, desc)), run(fill_flat)
    assert b[1] > 1000 and abs(a[1] - b[1]) <= 5.0 * np.sqrt(b[1]), (a[:2], b[:2])      # radiance photons
    assert np.isfinite(a[2]).all() and b[2].mean() > 1e-3
    assert abs(a[2].mean() - b[2].mean()) < 0.03 * b[2].mean(), (a[2].mean(), b[2].mean())


def test_instances_only_scene_and_tiny_objects(built):
    """edge cases of the two-level build: a world that holds nothing but instances, an object of ONE triangle (single-node
    tree), an object of one analytic sphere, and an identity instance that must return the bits of the plain mesh"""
    P, idx = S.heightfield_mesh(12, 12, None, seed=4)
    ident = S.xf_translate(0.0, 0.0, 0.0)
    tri_P = np.float32([[-0.5, 0.9, -0.5], [0.5, 0.9, -0.5], [0.0, 0.9, 0.6]])
    r = np.concatenate([S.pinhole_rays(64, 48, eye=(0.2, 3.0, 0.3), look=(0.0, 0.0, 0.0), up=(0, 0, 1)),
                        S.incoherent_rays(4000, [-1.2, -0.5, -1.2], [1.2, 1.5, 1.2], seed=6)])
    with b2rt.Context(0, seed=1) as c:
        o_mesh = c.object_begin(); c.add_mesh(P, idx); c.object_end()
        o_tri = c.object_begin(); c.add_mesh(tri_P, np.int32([[0, 1, 2]])); c.object_end()
        o_sph = c.object_begin(); c.add_sphere(S.xf_translate(0.0, 0.0, 0.0)[0], 0.2); c.object_end()
        k0 = c.add_instance(o_mesh, ident[0], w2o=ident[1])
        k1 = c.add_instance(o_tri, ident[0], w2o=ident[1])
        k2 = c.add_instance(o_sph, S.xf_translate(0.6, 1.3, 0.1)[0])
        info = c.build_accel()
        assert info.n_prims == 3
        bases = [c.instance_info(k)[0] for k in (k0, k1, k2)]
        assert bases == [3, 3 + len(idx), 3 + len(idx) + 1]
        h2 = c.trace_closest(r)
        o2 = c.trace_any(r)
    with b2rt.Context(0, seed=1) as f:
        f.add_mesh(P, idx)
        f.add_mesh(tri_P, np.int32([[0, 1, 2]]))
        f.add_sphere(S.xf_translate(0.6, 1.3, 0.1)[0], 0.2)
        f.build_accel()
        h1 = f.trace_closest(r)
        o1 = f.trace_any(r)
    mesh_or_tri = h1["prim"] <= len(idx)               # identity instances: the same bits as the plain shapes
    assert np.array_equal(h1["t"][mesh_or_tri].view(np.uint32), h2["t"][mesh_or_tri].view(np.uint32))
    assert np.array_equal(h1["prim"] + np.where(h1["prim"] >= 0, 3, 0), h2["prim"])
    assert np.allclose(h1["t"], h2["t"], rtol=1e-5) and np.array_equal(o1, o2)
    assert (h2["prim"] == 3 + len(idx)).sum() > 10 and (h2["prim"] == 3 + len(idx) + 1).sum() > 10
