"""Drop-in test (-m gpu): the reference's own pbrt-v2 front end, built with -DHAS_CUDA_RENDER against shim/ (the
cudaapi.h boundary) and libb2rt.so, renders a .pbrt scene file with `-c` (the reference's --force_cudarender switch,
pbrt-v2/src/main/pbrt.cpp:51) and is compared with the same binary's CPU path on the same file.
The binary is built in the build container (make -C shim; needs /root/reference) and travels to the GPU box."""
import os
import subprocess
import numpy as np
import pytest

from cuda_raytrace_b200 import scenes as S

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "shim", "_build", "pbrt_b2rt")
EYE, LOOK, UP, FOV = (0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40


def _run(tmp_path, name, integrator, renderer, spp, res=96, pixelfilter='"box"', sampler="lowdiscrepancy", desc=None,
         extra_world="", gpu_env=None):
    out = {}
    desc = desc or S.cornell_scene(boxes=True)
    for tag, flag in (("gpu", ["-c"]), ("cpu", [])):
        pf = str(tmp_path / ("%s_%s.pfm" % (name, tag)))
        scene = str(tmp_path / ("%s_%s.pbrt" % (name, tag)))
        with open(scene, "w") as f:
            f.write(S.to_pbrt(desc, EYE, LOOK, UP, FOV, res, res, spp, integrator, pf, renderer=renderer,
                              pixelfilter=pixelfilter, sampler=sampler, extra_world=extra_world))
        env = dict(os.environ, **(gpu_env or {})) if tag == "gpu" else dict(os.environ)
        r = subprocess.run([BIN, "--quiet"] + flag + [scene], capture_output=True, text=True, timeout=600, env=env)
        assert r.returncode == 0, r.stderr[-2000:]
        out[tag] = S.read_pfm(pf)
    return out["gpu"], out["cpu"]


def _block(img, b):
    h, w, _ = img.shape
    return img[: h // b * b, : w // b * b].reshape(h // b, b, w // b, b, 3).mean((1, 3))


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_simple_renderer_matches_cpu_directlighting(tmp_path):
    gpu, cpu = _run(tmp_path, "simple", 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]', "simple", spp=256)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.005, m
    assert m["mean_rel"] < 0.01, m


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_photon_renderer_matches_cpu_photonmap(tmp_path):
    prm = ('"integer causticphotons" [20000] "integer indirectphotons" [300000] "integer nused" [100] "float maxdist" [0.4] '
           '"integer maxphotondepth" [5] "integer maxspeculardepth" [5]')
    gpu, cpu = _run(tmp_path, "photon", 'SurfaceIntegrator "photonmap" "bool finalgather" ["false"] ' + prm,
                    'Renderer "photonmap" "bool finalgather" ["false"] ' + prm, spp=64)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.01, m
    assert m["mean_rel"] < 0.03, m


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_photon_renderer_final_gathering_default(tmp_path):
    """pbrt's DEFAULT photonmap mode (finalgather true, 32 gather samples, 10 degree cones) on both sides: the scene file
    says nothing about final gathering and the drop-in follows the integrator's defaults (photonmap.cpp:793-796)."""
    prm = ('"integer causticphotons" [20000] "integer indirectphotons" [300000] "integer nused" [100] "float maxdist" [0.4] '
           '"integer maxphotondepth" [5] "integer maxspeculardepth" [5]')
    gpu, cpu = _run(tmp_path, "photonfg", 'SurfaceIntegrator "photonmap" ' + prm, 'Renderer "photonmap" ' + prm, spp=64, res=64)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.03, m
    assert m["mean_rel"] < 0.06, m


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_honours_the_scene_pixel_filter(tmp_path):
    """`PixelFilter "mitchell"` in the scene file: the shim hands the film's filter to b2rt_film_set_filter and writes
    the reconstructed pixels unfiltered; the CPU path filters in ImageFilm::AddSample.  Compared per pixel (no block
    averaging) on a smoothed error metric, and against the box-filter render, which must differ."""
    integ = 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]'
    flt = '"mitchell" "float xwidth" [2] "float ywidth" [2]'
    gpu, cpu = _run(tmp_path, "mitchell", integ, "simple", spp=256, res=64, pixelfilter=flt)
    gbox, _ = _run(tmp_path, "box", integ, "simple", spp=256, res=64)
    m = S.image_metrics(_block(gpu, 4), _block(cpu, 4))
    assert abs(m["bias_rel"]) < 0.005, m
    assert m["mean_rel"] < 0.015, m
    # the Mitchell reconstruction is visibly different from the box one at edges (negative lobes, wider support)
    edge = np.abs(gpu - gbox).max()
    assert edge > 0.02 * gbox.max(), edge
    # and per pixel the GPU image is closer to pbrt's Mitchell image than the box image is
    assert np.abs(gpu - cpu).mean() < np.abs(gbox - cpu).mean()


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
@pytest.mark.parametrize("sampler", ["stratified", "random"])
def test_dropin_follows_the_scene_sampler(tmp_path, sampler):
    """`Sampler "stratified"` (16 x 16 strata) and `Sampler "random"`: the shim passes the pattern on with
    b2rt_sampler_set; both backends converge to the same image."""
    gpu, cpu = _run(tmp_path, sampler, 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]', "simple", spp=256,
                    res=64, sampler=sampler)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.005, m
    assert m["mean_rel"] < 0.012, m


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_mesh_with_vertex_normals_and_uv(tmp_path):
    """a `Shape "trianglemesh"` with `"normal N"` and `"float uv"` under a non-uniform object transform.
    The shim reads TriangleMesh::n / uvs through the fork's friend declaration, applies the object-to-world NORMAL
    transform (inverse transpose) and hands both to b2rt_add_mesh_ex; pbrt's CPU path interpolates the shading normal in
    Triangle::GetShadingGeometry.  A coarse ball (12 x 6) makes the difference between smooth and faceted shading large:
    the GPU image must match pbrt's smooth image and must NOT match pbrt's image of the same mesh without normals."""
    integ = 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]'
    P, idx, N, uv = S.sphere_mesh(12, 6, 1.0, (0.0, 0.0, 0.0))
    o2w = [0.45, 0, 0, 0.0, 0, 0.6, 0, 0.6, 0, 0, 0.45, 0.1]        # Translate(0, .6, .1) Scale(.45, .6, .45): an ellipsoid
    desc = S.cornell_scene(spheres=False)
    desc["shapes"].append(("mesh", P, idx, 0, N, uv, o2w))
    gpu, cpu = _run(tmp_path, "smooth", integ, "simple", spp=256, res=96, desc=desc)
    flat = S.cornell_scene(spheres=False)
    flat["shapes"].append(("mesh", P, idx, 0, None, None, o2w))
    _, cpu_flat = _run(tmp_path, "flat", integ, "simple", spp=256, res=96, desc=flat)
    m = S.image_metrics(_block(gpu, 4), _block(cpu, 4))
    assert abs(m["bias_rel"]) < 0.005 and m["mean_rel"] < 0.015, m
    ball = (slice(40, 88), slice(28, 68))                            # the ellipsoid's pixels
    mb = S.image_metrics(_block(gpu[ball], 4), _block(cpu[ball], 4))
    mf = S.image_metrics(_block(gpu[ball], 4), _block(cpu_flat[ball], 4))
    assert mb["mean_rel"] < 0.015 and mf["mean_rel"] > 2 * mb["mean_rel"], (mb, mf)


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_integer_gpus_gives_the_same_image(tmp_path, monkeypatch):
    """`Renderer "photonmap" "integer gpus" [3]`: the shim grows its GPU group (b2rt_multi_resize copies the parsed scene),
    shoots the photons sharded, renders image tiles and composites on the first GPU -- the PFM must be byte-identical to
    the one-GPU run.  On a box with fewer GPUs the group is built from contexts on one device (B2RT_GPUS_ON_ONE_DEVICE)."""
    import torch
    if torch.cuda.device_count() < 3:
        monkeypatch.setenv("B2RT_GPUS_ON_ONE_DEVICE", "1")
    prm = ('"integer causticphotons" [5000] "integer indirectphotons" [100000] "integer nused" [50] "float maxdist" [0.3] '
           '"bool finalgather" ["false"]')
    imgs = []
    for g in (1, 3):
        pf = str(tmp_path / ("gpus%d.pfm" % g))
        scene = str(tmp_path / ("gpus%d.pbrt" % g))
        with open(scene, "w") as f:
            f.write(S.to_pbrt(S.cornell_scene(), EYE, LOOK, UP, FOV, 160, 100, 4, 'SurfaceIntegrator "photonmap" ' + prm, pf,
                              renderer='Renderer "photonmap" "integer gpus" [%d] %s' % (g, prm)))
        r = subprocess.run([BIN, "--quiet", "-c", scene], capture_output=True, text=True, timeout=600, env=dict(os.environ))
        assert r.returncode == 0, r.stderr[-2000:]
        imgs.append(S.read_pfm(pf))
    assert imgs[0].max() > 0 and np.array_equal(imgs[0], imgs[1])


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
@pytest.mark.parametrize("light", ["spot", "distant"])
def test_dropin_spot_and_distant_lights(tmp_path, light):
    """`LightSource "spot"` / `"distant"` in the scene file (the reference's GPU path skips both with a warning): direct
    lighting and photon map against pbrt's CPU renderer on the same file."""
    desc = S.cornell_scene(boxes=True, light=light)
    gpu, cpu = _run(tmp_path, light + "_direct", 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]', "simple", spp=64,
                    res=64, desc=desc)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.005 and m["mean_rel"] < 0.012, m
    prm = ('"integer causticphotons" [10000] "integer indirectphotons" [200000] "integer nused" [100] "float maxdist" [0.4] '
           '"bool finalgather" ["false"]')
    gpu, cpu = _run(tmp_path, light + "_photon", 'SurfaceIntegrator "photonmap" ' + prm, 'Renderer "photonmap" ' + prm, spp=16,
                    res=64, desc=desc)
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.015 and m["mean_rel"] < 0.04, m


def _instanced_world():
    """an ObjectBegin block (a smooth ball mesh with normals, an analytic sphere and a disk, each with its own material) and
    five ObjectInstance lines under rotations, non-uniform scales and a mirroring transform"""
    P, idx, N, uv = S.sphere_mesh(12, 6, 0.5, (0.0, 0.0, 0.0))
    L = ['ObjectBegin "thing"',
         '  AttributeBegin', '    Material "matte" "rgb Kd" [0.7 0.3 0.2]',
         '    Shape "trianglemesh" "integer indices" [%s] "point P" [%s] "normal N" [%s]' % (
             " ".join(str(int(v)) for v in idx.reshape(-1)), " ".join("%.9g" % v for v in P.reshape(-1)),
             " ".join("%.9g" % v for v in N.reshape(-1))),
         '  AttributeEnd',
         '  AttributeBegin', '    Material "mirror" "rgb Kr" [0.9 0.9 0.9]', '    Translate 0 0.75 0',
         '    Shape "sphere" "float radius" [0.25]', '  AttributeEnd',
         '  AttributeBegin', '    Material "matte" "rgb Kd" [0.2 0.6 0.3]', '    Translate 0 -0.5 0', '    Rotate -90 1 0 0',
         '    Shape "disk" "float radius" [0.7]', '  AttributeEnd',
         'ObjectEnd']
    for (t, r, sc) in (((-0.55, 0.45, -0.3), (20, 0, 1, 0), (0.5, 0.5, 0.5)), ((0.5, 0.35, -0.1), (35, 1, 1, 0), (0.45, 0.3, 0.45)),
                       ((0.0, 0.4, 0.45), (-30, 0, 1, 1), (-0.35, 0.35, 0.35)), ((-0.45, 1.3, 0.2), (80, 1, 0, 0), (0.3, 0.5, 0.3)),
                       ((0.5, 1.35, -0.4), (0, 0, 1, 0), (0.4, 0.4, 0.4))):
        L += ['AttributeBegin', '  Translate %g %g %g' % t, '  Rotate %g %g %g %g' % r, '  Scale %g %g %g' % sc,
              '  ObjectInstance "thing"', 'AttributeEnd']
    return "\n".join(L)


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_object_instances_two_level(tmp_path):
    """`ObjectBegin` / `ObjectInstance` in the scene file: the shim makes ONE b2rt object and five instances of it (two-level
    BVH; the reference hangs them under OptiX Transform nodes, cudarender.cpp:38-103) -- against pbrt's CPU image of the
    same file (TransformedPrimitive), and against the same run with B2RT_FLATTEN_INSTANCES=1 (every instance written out)."""
    integ = 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]'
    desc = S.cornell_scene(spheres=False)
    world = _instanced_world()
    gpu, cpu = _run(tmp_path, "inst", integ, "simple", spp=256, res=96, desc=desc, extra_world=world)
    m = S.image_metrics(_block(gpu, 4), _block(cpu, 4))
    assert abs(m["bias_rel"]) < 0.005 and m["mean_rel"] < 0.015, m
    flat, _ = _run(tmp_path, "instflat", integ, "simple", spp=256, res=96, desc=desc, extra_world=world,
                   gpu_env={"B2RT_FLATTEN_INSTANCES": "1"})
    d = np.abs(gpu - flat).sum(-1) > 1e-3 * (1.0 + flat.sum(-1))
    assert d.mean() < 0.02, d.mean()
    empty, _ = _run(tmp_path, "noinst", integ, "simple", spp=16, res=96, desc=desc)
    assert np.abs(gpu - empty).mean() > 0.01 * empty.mean()          # the instances are in the picture


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_object_instances_photonmap(tmp_path):
    """the same instanced file through the photon-mapping renderer (photon pass + final gathering trace the two-level tree)"""
    prm = ('"integer causticphotons" [10000] "integer indirectphotons" [200000] "integer nused" [100] "float maxdist" [0.4] '
           '"bool finalgather" ["false"]')
    gpu, cpu = _run(tmp_path, "instpm", 'SurfaceIntegrator "photonmap" ' + prm, 'Renderer "photonmap" ' + prm, spp=64,
                    desc=S.cornell_scene(spheres=False), extra_world=_instanced_world())
    m = S.image_metrics(_block(gpu, 8), _block(cpu, 8))
    assert abs(m["bias_rel"]) < 0.01 and m["mean_rel"] < 0.03, m


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_partial_sphere(tmp_path):
    """`Shape "sphere" "float zmax" "float phimax"` (a bowl with a slice cut out, open towards the camera) and a 250-degree ring
    (`Shape "disk" "float phimax"`).  The reference's device code renders the full shapes; the drop-in passes zmin / zmax /
    phimax on (b2rt_add_sphere_ex, b2rt_add_disk_ex) and must match pbrt's CPU image, inside of the bowl included."""
    integ = 'SurfaceIntegrator "directlighting" "integer maxdepth" [5]'
    bowl = "\n".join(['AttributeBegin', '  Material "matte" "rgb Kd" [0.7 0.5 0.2]', '  Translate 0 0.7 0', '  Rotate -60 1 0 0',
                      '  Shape "sphere" "float radius" [0.6] "float zmax" [0.15] "float phimax" [300]', 'AttributeEnd',
                      'AttributeBegin', '  Material "matte" "rgb Kd" [0.2 0.4 0.7]', '  Translate -0.55 1.4 -0.2', '  Rotate 60 1 0 0',
                      '  Shape "disk" "float radius" [0.35] "float innerradius" [0.1] "float phimax" [250]', 'AttributeEnd'])
    desc = S.cornell_scene(spheres=False)
    gpu, cpu = _run(tmp_path, "bowl", integ, "simple", spp=256, res=96, desc=desc, extra_world=bowl)
    m = S.image_metrics(_block(gpu, 4), _block(cpu, 4))
    assert abs(m["bias_rel"]) < 0.005 and m["mean_rel"] < 0.015, m
    full = bowl.replace(' "float zmax" [0.15] "float phimax" [300]', "")
    _, cpu_full = _run(tmp_path, "ball", integ, "simple", spp=64, res=96, desc=desc, extra_world=full)
    mf = S.image_metrics(_block(gpu, 4), _block(cpu_full, 4))
    assert mf["mean_rel"] > 3 * m["mean_rel"], (m, mf)                 # and it is not the full sphere


@pytest.mark.skipif(not os.path.exists(BIN), reason="shim/_build/pbrt_b2rt not built (needs /root/reference)")
def test_dropin_scene_cache(tmp_path):
    """`Renderer "photonmap" "string savescene"` writes the parsed scene (instances, materials, area light) to a cache;
    a stub file with no shapes and `"string loadscene"` renders it again without the text parser seeing a triangle: same PFM"""
    prm = '"integer causticphotons" [2000] "integer indirectphotons" [50000] "integer nused" [50] "float maxdist" [0.4] "bool finalgather" ["false"]'
    cache = str(tmp_path / "scene.b2rt")
    out = []
    for tag, desc, extra, rp in (("full", S.cornell_scene(spheres=True), _instanced_world(), '"string savescene" ["%s"]' % cache),
                                 ("stub", {"materials": [], "shapes": [], "lights": []}, "", '"string loadscene" ["%s"]' % cache)):
        pf = str(tmp_path / (tag + ".pfm"))
        scene = str(tmp_path / (tag + ".pbrt"))
        with open(scene, "w") as f:
            f.write(S.to_pbrt(desc, EYE, LOOK, UP, FOV, 96, 96, 4, 'SurfaceIntegrator "photonmap" ' + prm, pf,
                              renderer='Renderer "photonmap" %s %s' % (rp, prm), extra_world=extra))
        r = subprocess.run([BIN, "--quiet", "-c", scene], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        out.append(S.read_pfm(pf))
    assert os.path.getsize(cache) > 4000 and os.path.getsize(str(tmp_path / "stub.pbrt")) < 2000
    assert out[0].max() > 0 and np.array_equal(out[0], out[1])
