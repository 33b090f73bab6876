"""GPU parity (through the C ABI): the film reconstruction kernel (csrc/film.cu) against oracle/oracle_film.c - itself
bit-identical to pbrt's ImageFilm (tests/test_oracle_film_cpu.py) - and against the committed images of the reference's
own film (tests/golden/film.npz); the filtered camera pass: band tiling, constant-radiance and box-filter identities."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
from cuda_raytrace_b200 import b2rt, scenes as S
import film_cases as F

pytestmark = pytest.mark.gpu


def _flt(name, kw):
    return b2rt.pbrt_filter(name, **kw), O.filter_args(name, **kw)


@pytest.mark.parametrize("case", range(len(F.CASES)))
def test_film_kernel_matches_oracle_and_pbrt_golden_bit_for_bit(case):
    name, kw, xres, yres, spp, window = F.CASES[case]
    gf, of = _flt(name, kw)
    cells, xy, rgb, px = F.make_samples(of, xres, yres, spp, window, seed=100 + case)
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, spp, xy, rgb, px)
    ref = O.film(of, cells, spp, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))
    assert np.array_equal(img.view(np.uint32), load_golden("film")["film_%d" % case].view(np.uint32))


@pytest.mark.parametrize("variant", ["1", "0"])
def test_film_kernel_at_frame_size_against_oracle(variant, monkeypatch):
    # 1920x1080, 2 spp, mitchell: 4.2 M samples; the oracle finishes in seconds.  Variant 1 = shared-memory tiles
    # (default), 0 = one thread per pixel; B2RT_FILM_VARIANT is read when the context is created
    monkeypatch.setenv("B2RT_FILM_VARIANT", variant)
    gf, of = _flt("mitchell", {})
    cells, xy, rgb, px = F.make_samples(of, 1920, 1080, 2, None, seed=9)
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, 2, xy, rgb, px)
    ref = O.film(of, cells, 2, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("case", range(len(F.CASES)))
def test_film_per_pixel_variant_matches_golden(case, monkeypatch):
    monkeypatch.setenv("B2RT_FILM_VARIANT", "0")
    name, kw, xres, yres, spp, window = F.CASES[case]
    gf, of = _flt(name, kw)
    cells, xy, rgb, px = F.make_samples(of, xres, yres, spp, window, seed=100 + case)
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, spp, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), load_golden("film")["film_%d" % case].view(np.uint32))


def test_film_wide_filter_and_many_samples_per_pixel():
    # sinc of width 6.5 (halo 7) and 40 spp: a staged cell row is 47 x 40 samples = 37 KB of shared memory; odd sizes
    gf = b2rt.pbrt_filter("sinc", xwidth=6.5, ywidth=3.0, tau=2.0)
    of = O.filter_args("sinc", xwidth=6.5, ywidth=3.0, tau=2.0)
    cells, xy, rgb, px = F.make_samples(of, 77, 45, 40, (5, 3, 70, 41), seed=21)
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, 40, xy, rgb, px)
    ref = O.film(of, cells, 40, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


def test_film_argument_errors():
    with b2rt.Context(0) as c:
        xy, rgb = np.zeros((4, 2), np.float32), np.zeros((4, 3), np.float32)
        with pytest.raises(b2rt.B2rtError):
            c.film_reconstruct(b2rt.Filter(7, 1.0, 1.0, 0, 0), (0, 0, 2, 2), 1, xy, rgb, (0, 0, 2, 2))
        with pytest.raises(b2rt.B2rtError):
            c.film_reconstruct(b2rt.Filter(0, 0.0, 1.0, 0, 0), (0, 0, 2, 2), 1, xy, rgb, (0, 0, 2, 2))
        with pytest.raises(b2rt.B2rtError):
            c.set_filter(b2rt.Filter(3, 40.0, 2.0, 0, 0))


EYE, LOOK, UP, FOV = (0, 1, 3.4), (0, 1, 0), (0, 1, 0), 40


@pytest.fixture(scope="module")
def rctx():
    with b2rt.Context(0, seed=5) as c:
        S.fill_scene(c, S.cornell_scene(spheres=True))
        c.build_accel()
        c.photons_shoot(2000, 20000, batch_paths=32768)
        c.grid_build(0)
        c.grid_build(1)
        yield c


def test_filtered_frame_bands_are_bit_identical_and_filter_resets(rctx):
    cam = b2rt.pbrt_camera(EYE, LOOK, UP, FOV, 64, 48)
    kw = dict(spp=2, nused=20, max_dist=0.2)
    box = rctx.render_frame(cam, **kw)
    rctx.set_filter(b2rt.pbrt_filter("mitchell"))
    full = rctx.render_frame(cam, **kw)
    again = rctx.render_frame(cam, **kw)
    assert np.array_equal(full.view(np.uint32), again.view(np.uint32))
    bands = [rctx.render_frame(cam, window=(0, 64, y0, y1), **kw) for y0, y1 in ((0, 13), (13, 30), (30, 48))]
    assert np.array_equal(np.concatenate(bands, 0).view(np.uint32), full.view(np.uint32))
    cols = [rctx.render_frame(cam, window=(x0, x1, 0, 48), **kw) for x0, x1 in ((0, 31), (31, 64))]
    assert np.array_equal(np.concatenate(cols, 1).view(np.uint32), full.view(np.uint32))
    assert not np.array_equal(full, box)
    # a wider reconstruction filter redistributes radiance between neighbouring pixels, it does not change the total much
    assert abs(full.sum() - box.sum()) <= 0.03 * box.sum()
    rctx.set_filter(None)
    assert np.array_equal(rctx.render_frame(cam, **kw).view(np.uint32), box.view(np.uint32))


def test_wide_box_filter_frame_equals_film_of_neighbouring_box_frames(rctx):
    # box filter of width 1.5: every pixel is the mean of all samples within 1.5 pixels; in the interior that is the
    # 3 x 3 block mean plus the samples of the next ring whose jitter falls inside - bounded by the block extremes
    cam = b2rt.pbrt_camera(EYE, LOOK, UP, FOV, 48, 48)
    kw = dict(spp=4, caustic=False, indirect=False)
    box = rctx.render_frame(cam, **kw).astype(np.float64)
    rctx.set_filter(b2rt.pbrt_filter("box", xwidth=1.0, ywidth=1.0))
    wide = rctx.render_frame(cam, **kw).astype(np.float64)
    rctx.set_filter(None)
    # width 1.0 box: pixel (x, y) averages the samples with |X - 0.5 - x| <= 1: its own cell and parts of the 8 around
    lo = np.full_like(box, np.inf); hi = np.full_like(box, -np.inf)
    pad = np.pad(box, ((1, 1), (1, 1), (0, 0)), mode="edge")
    for dy in range(3):
        for dx in range(3):
            lo = np.minimum(lo, pad[dy:dy + 48, dx:dx + 48]); hi = np.maximum(hi, pad[dy:dy + 48, dx:dx + 48])
    inner = (slice(2, 46), slice(2, 46))
    # per-pixel means of the neighbours bound the sample values only on smooth regions; check the image-wide mean and
    # that most pixels stay inside the neighbourhood range
    assert abs(wide[inner].mean() - box[inner].mean()) <= 0.02 * box[inner].mean()
    inside = (wide >= lo - 1e-6) & (wide <= hi + 1e-6)
    assert inside[inner].mean() > 0.8


@pytest.mark.parametrize("name,kw", [("box", {}), ("box", {"xwidth": 1.0, "ywidth": 0.75}), ("triangle", {}),
                                     ("mitchell", {"xwidth": 1.5, "ywidth": 2.25})])
@pytest.mark.parametrize("variant", ["1", "0"])
def test_film_border_samples_match_oracle(name, kw, variant, monkeypatch):
    # samples exactly on cell borders / pixel centres (the oracle is pinned to pbrt on the same set, CPU suite)
    monkeypatch.setenv("B2RT_FILM_VARIANT", variant)
    gf, of = _flt(name, kw)
    cells, xy, rgb, px = F.make_lattice_samples(of, 13, 9, (1, 1, 12, 8))
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, 16, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), O.film(of, cells, 16, xy, rgb, px).view(np.uint32))


# ---- image-plane sample patterns -----------------------------------------------------------------------------------------
@pytest.mark.parametrize("sampler,nx,ny,spp", [("random", 1, 1, 5), ("stratified", 3, 2, 6), ("stratified", 4, 4, 16),
                                               ("lowdiscrepancy", 1, 1, 8), ("lowdiscrepancy", 1, 1, 64)])
def test_sample_positions_match_oracle_bit_for_bit(sampler, nx, ny, spp):
    for seed, xres, yres, win in ((99, 12, 8, (-2, -1, 14, 9)), (2 ** 40 + 7, 640, 360, (100, 50, 300, 120))):
        with b2rt.Context(0, seed=seed) as c:
            c.set_sampler(sampler, nx, ny)
            xy = c.sample_positions(xres, yres, win, spp)
        ref = O.sample_positions(sampler, seed, xres, yres, win, spp, nx, ny)
        assert np.array_equal(xy.view(np.uint32), ref.view(np.uint32))


def test_sampler_errors_and_frames(rctx):
    cam = b2rt.pbrt_camera(EYE, LOOK, UP, FOV, 48, 32)
    kw = dict(spp=4, caustic=False, indirect=False)
    base = rctx.render_frame(cam, **kw)
    with pytest.raises(b2rt.B2rtError):
        rctx.set_sampler(7)
    rctx.set_sampler("stratified", 3, 2)
    with pytest.raises(b2rt.B2rtError):
        rctx.render_frame(cam, **kw)                                # 3 x 2 strata != 4 spp
    imgs = {}
    for name, nx, ny in (("stratified", 2, 2), ("lowdiscrepancy", 1, 1)):
        rctx.set_sampler(name, nx, ny)
        imgs[name] = rctx.render_frame(cam, **kw)
        bands = [rctx.render_frame(cam, window=(0, 48, y0, y1), **kw) for y0, y1 in ((0, 11), (11, 32))]
        assert np.array_equal(np.concatenate(bands, 0).view(np.uint32), imgs[name].view(np.uint32))
        assert not np.array_equal(imgs[name], base)
        assert abs(imgs[name].mean() - base.mean()) <= 0.1 * base.mean()      # 48 x 32 x 4 spp: Monte-Carlo noise
    rctx.set_sampler("random")
    assert np.array_equal(rctx.render_frame(cam, **kw).view(np.uint32), base.view(np.uint32))


def test_pattern_samplers_converge_faster_than_random(rctx):
    """Direct lighting of the area light, 4 spp: with the stratified / low-discrepancy samplers the pixel positions AND
    the light / BSDF samples of the camera hit come from one pattern per pixel, so the error against a converged image
    must be clearly below the independent-uniform sampler's (same estimator, same expectation)."""
    cam = b2rt.pbrt_camera(EYE, LOOK, UP, FOV, 96, 64)
    kw = dict(caustic=False, indirect=False)
    rctx.set_sampler("lowdiscrepancy")
    ref = rctx.render_frame(cam, spp=2048, **kw).astype(np.float64)
    err = {}
    for name, nx, ny in (("random", 1, 1), ("stratified", 2, 2), ("lowdiscrepancy", 1, 1)):
        rctx.set_sampler(name, nx, ny)
        img = rctx.render_frame(cam, spp=4, **kw).astype(np.float64)
        err[name] = float(np.sqrt(((img - ref) ** 2).mean()))
        assert abs(img.mean() - ref.mean()) <= 0.02 * ref.mean(), (name, img.mean(), ref.mean())
    rctx.set_sampler("random")
    assert err["lowdiscrepancy"] < 0.85 * err["random"], err
    assert err["stratified"] < 0.95 * err["random"], err


def test_film_full_size_properties():
    """BASELINE Cfg4 frame size (3840 x 2160, 8 spp, Mitchell): size-independent properties of the reconstruction -
    a constant radiance field comes back as that constant, two runs and two band tilings give the same bits, and a
    random sample of pixels equals the oracle's value for the same samples."""
    gf, of = _flt("mitchell", {})
    xres, yres, spp = 3840, 2160, 8
    h = F.halo(of[1])
    cells = (-h, -h, xres + h, yres + h)
    cw, ch = cells[2] - cells[0], cells[3] - cells[1]
    rng = np.random.default_rng(31)
    xy = rng.random((ch, cw, spp, 2), dtype=np.float32)
    xy[..., 0] += np.arange(cells[0], cells[2], dtype=np.float32)[None, :, None]
    xy[..., 1] += np.arange(cells[1], cells[3], dtype=np.float32)[:, None, None]
    xy = xy.reshape(-1, 2)
    rgb = rng.random((len(xy), 3), dtype=np.float32)
    with b2rt.Context(0) as c:
        img = c.film_reconstruct(gf, cells, spp, xy, rgb, (0, 0, xres, yres))
        again = c.film_reconstruct(gf, cells, spp, xy, rgb, (0, 0, xres, yres))
        assert np.array_equal(img.view(np.uint32), again.view(np.uint32))
        # a band of the frame from only the cell rows that reach it
        y0, y1 = 1000, 1100
        rows = slice((y0 - h - cells[1]) * cw * spp, (y1 + h - cells[1]) * cw * spp)
        band = c.film_reconstruct(gf, (cells[0], y0 - h, cells[2], y1 + h), spp, xy[rows], rgb[rows], (0, y0, xres, y1))
        assert np.array_equal(band.view(np.uint32), img[y0:y1].view(np.uint32))
        const = c.film_reconstruct(gf, cells, spp, xy, np.full_like(rgb, 0.75), (0, 0, xres, yres))
        assert np.abs(const - np.float32(0.75)).max() <= 2e-5
    # oracle on a 64 x 48 window of the same samples
    wx0, wy0, wx1, wy1 = 1900, 1050, 1964, 1098
    sel = []
    for cy in range(wy0 - h, wy1 + h):
        a = ((cy - cells[1]) * cw + (wx0 - h - cells[0])) * spp
        sel.append(np.arange(a, a + (wx1 - wx0 + 2 * h) * spp))
    sel = np.concatenate(sel)
    ref = O.film(of, (wx0 - h, wy0 - h, wx1 + h, wy1 + h), spp, xy[sel], rgb[sel], (wx0, wy0, wx1, wy1))
    assert np.array_equal(ref.view(np.uint32), img[wy0:wy1, wx0:wx1].view(np.uint32))
