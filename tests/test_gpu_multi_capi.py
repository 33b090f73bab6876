"""GPU tests (-m gpu) of the multi-GPU layer of the C ABI (csrc/multi.cu, b2rt_multi_*): sharded photon passes, replicated
grids, tiled camera pass.  Bar: every result byte-identical to the single-context calls, for every GPU count and every
exchange mode.  On a one-GPU box the group is made of several contexts on the SAME device (the sharding, the offsets, the
multi-destination compaction and the tile lists are exercised all the same; the stores just do not cross NVLink); with
two or more GPUs the real devices are used as well."""
import hashlib
import os

import numpy as np
import pytest

from cuda_raytrace_b200 import b2rt, scenes as S

pytestmark = pytest.mark.gpu
N_PATHS = 300_000
CAM = dict(eye=(0, 1, 3.4), look=(0, 1, 0), up=(0, 1, 0), fov=40)


def _groups():
    out = [("1ctx", [0]), ("3ctx_same_gpu", [0, 0, 0])]
    n = b2rt.load().b2rt_device_count()
    if n >= 2:
        out.append(("2gpus", [0, 1]))
    if n >= 4:
        out.append(("%dgpus" % n, list(range(n))))
    return out


@pytest.fixture(scope="module")
def single(built):
    """single-context reference results"""
    desc = S.cornell_scene(wide=1.78)
    cam = b2rt.pbrt_camera(CAM["eye"], CAM["look"], CAM["up"], CAM["fov"], 200, 120)
    out = {}
    with b2rt.Context(0, seed=3) as c:
        S.fill_scene(c, desc)
        c.build_accel()
        c.photons_trace(0, N_PATHS)
        out["maps"] = [c.photons_download(m).tobytes() for m in (0, 1)]
        c.grid_build(0); c.grid_build(1)
        out["box"] = c.render_frame(cam, spp=2, nused=40, max_dist=0.15)
        out["crop"] = c.render_frame(cam, spp=2, nused=40, max_dist=0.15, window=(30, 171, 17, 101))
        c.set_filter(b2rt.pbrt_filter("mitchell")); c.set_sampler("lowdiscrepancy")
        out["mitchell"] = c.render_frame(cam, spp=4, nused=40, max_dist=0.15)
        c.set_filter(None); c.set_sampler("random")
        c.photons_trace_fg(0, N_PATHS)
        out["fgmaps"] = [c.photons_download(m).tobytes() for m in (2, 3)]
        c.grid_build(2)
        c.radiance_compute(40, 0.15)
        out["rad"] = c.photons_download(3).tobytes()
        c.set_final_gather(True, 8, 10.0)
        out["fg"] = c.render_frame(cam, spp=1, nused=40, max_dist=0.15)
        (pc, pi), (nc, ni) = c.photons_shoot(300, 40000, batch_paths=16384)
        out["shoot"] = ((pc, pi), (nc, ni), hashlib.sha256(c.photons_download(1).tobytes()).hexdigest())
    return desc, cam, out


@pytest.mark.parametrize("mode", ["fused", "push", "copy", "nccl"])
@pytest.mark.parametrize("group", _groups(), ids=[g[0] for g in _groups()])
def test_multi_matches_single_context(single, group, mode, monkeypatch):
    desc, cam, ref = single
    devices = group[1]
    if mode == "nccl" and len(set(devices)) != len(devices):
        pytest.skip("NCCL needs distinct devices")
    monkeypatch.setenv("B2RT_MGPU_EXCHANGE", mode)
    with b2rt.MultiContext(len(devices), seed=3, devices=devices) as m:
        info = m.info()
        assert info["gpus"] == len(devices)
        m.fill(lambda c: S.fill_scene(c, desc))
        m.build_accel()
        nc, ni = m.photons_trace(0, N_PATHS)
        for c in m.ctxs:                                                 # EVERY context holds the complete maps
            assert [c.photons_download(k).tobytes() for k in (0, 1)] == ref["maps"]
            assert c.photons_count(1) == (ni, N_PATHS)
        m.grid_build(0); m.grid_build(1)
        kw = dict(nused=40, max_dist=0.15)
        assert np.array_equal(m.render_frame(cam, spp=2, **kw), ref["box"])                 # image tiles
        assert np.array_equal(m.render_frame(cam, spp=2, window=(30, 171, 17, 101), **kw), ref["crop"])
        m.set_filter(b2rt.pbrt_filter("mitchell")); m.set_sampler("lowdiscrepancy")
        assert np.array_equal(m.render_frame(cam, spp=4, **kw), ref["mitchell"])            # row bands with halo cells
        m.set_filter(None); m.set_sampler("random")
        m.photons_trace_fg(0, N_PATHS)
        for c in m.ctxs:
            assert [c.photons_download(k).tobytes() for k in (2, 3)] == ref["fgmaps"]
        m.grid_build(2)
        m.radiance_compute(40, 0.15)
        assert all(c.photons_download(3).tobytes() == ref["rad"] for c in m.ctxs)
        m.set_final_gather(True, 8, 10.0)
        assert np.array_equal(m.render_frame(cam, spp=1, **kw), ref["fg"])
        m.set_final_gather(False)
        (pc, pi), (nc, ni) = m.photons_shoot(300, 40000, batch_paths=16384)                 # pbrt's stopping rule, sharded
        assert ((pc, pi), (nc, ni), hashlib.sha256(m.ctxs[-1].photons_download(1).tobytes()).hexdigest()) == ref["shoot"]
        st = m.stats()
        assert st["passes"] >= 3 and st["exchange_bytes"] > 0


def test_multi_ragged_and_empty_slices(built):
    """more contexts than paths, zero paths, a scene whose caustic map stays empty"""
    desc = S.cornell_scene(spheres=False, boxes=True)
    with b2rt.Context(0, seed=9) as c:
        S.fill_scene(c, desc); c.build_accel()
        c.photons_trace(5, 3)
        want = [c.photons_download(k).tobytes() for k in (0, 1)]
    with b2rt.MultiContext(4, seed=9, devices=[0, 0, 0, 0]) as m:
        m.fill(lambda c: S.fill_scene(c, desc)); m.build_accel()
        assert m.photons_trace(0, 0) == (0, 0)
        m.photons_trace(5, 3)
        assert [m.ctxs[3].photons_download(k).tobytes() for k in (0, 1)] == want
        with pytest.raises(b2rt.B2rtError):
            m.photons_trace(0, 10, max_depth=0)


def test_multi_resize_copies_the_scene(single):
    """the drop-in's order of events: the scene goes into ONE context while the file is parsed, the GPU count arrives with
    the Renderer line; b2rt_multi_resize clones the staged scene and settings into the new contexts"""
    desc, cam, ref = single
    with b2rt.MultiContext(1, seed=3, devices=[0]) as m:
        S.fill_scene(m.ctxs[0], desc)
        m.resize(3, devices=[0, 0, 0])
        m.build_accel()
        m.photons_trace(0, N_PATHS)
        assert [m.ctxs[2].photons_download(k).tobytes() for k in (0, 1)] == ref["maps"]
        m.grid_build(0); m.grid_build(1)
        assert np.array_equal(m.render_frame(cam, spp=2, nused=40, max_dist=0.15), ref["box"])


def test_multi_resize_copies_objects_and_instances(built):
    """the same order of events for a scene with ObjectInstances: the objects (child contexts) and the instance list are
    cloned too, every device builds the same two-level tree, and maps and frame equal the single-context results"""
    desc = S.instanced_scene(n_inst=6, seed=13)
    cam = b2rt.pbrt_camera((0.5, 3.0, 7.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), 50.0, 128, 96)

    def fill(c):
        S.fill_instanced(c, desc)
        c.add_light(0, (60.0, 60.0, 60.0), p=(0.5, 4.5, 1.0))

    with b2rt.Context(0, seed=7) as c:
        fill(c)
        c.build_accel()
        c.photons_trace(0, 100_000)
        want = [c.photons_download(k).tobytes() for k in (0, 1)]
        c.grid_build(1)
        img = c.render_frame(cam, spp=2, nused=30, max_dist=0.5)
    with b2rt.MultiContext(1, seed=7, devices=[0]) as m:
        fill(m.ctxs[0])
        m.resize(3, devices=[0, 0, 0])
        m.build_accel()
        m.photons_trace(0, 100_000)
        assert [m.ctxs[2].photons_download(k).tobytes() for k in (0, 1)] == want
        m.grid_build(1)
        assert np.array_equal(m.render_frame(cam, spp=2, nused=30, max_dist=0.5), img)
        assert img.mean() > 1e-3


def test_multi_frame_straight_into_pinned_host_memory(single, monkeypatch):
    """a whole frame asked for in page-locked host memory is stored there by the GPUs themselves (no collection on GPU 0,
    no final download); the bytes are those of the staged path (B2RT_MGPU_HOSTFRAME=0) and of a pageable output buffer"""
    desc, cam, ref = single
    with b2rt.MultiContext(3, seed=3, devices=[0, 0, 0]) as m:
        m.fill(lambda c: S.fill_scene(c, desc))
        m.build_accel()
        m.photons_trace(0, N_PATHS)
        m.grid_build(0); m.grid_build(1)
        pin = m.ctxs[0].pinned_empty((cam.yres, cam.xres, 3))
        pin[:] = -1.0
        monkeypatch.setenv("B2RT_MGPU_HOSTFRAME", "1")            # default: on from 4 GPUs
        got = m.render_frame(cam, spp=2, nused=40, max_dist=0.15, out=pin)
        assert got is pin and np.array_equal(pin, ref["box"])
        monkeypatch.setenv("B2RT_MGPU_HOSTFRAME", "0")
        pin[:] = -1.0
        m.render_frame(cam, spp=2, nused=40, max_dist=0.15, out=pin)
        assert np.array_equal(pin, ref["box"])
        assert np.array_equal(m.render_frame(cam, spp=2, nused=40, max_dist=0.15), ref["box"])
