"""CPU tests (no GPU): oracle/oracle_film.c, the restatement of pbrt-v2's ImageFilm (film/image.cpp:41-137,178-213) and
its pixel filters (filters/*.cpp), pinned against
  * tests/golden/film.npz - images the reference's own ImageFilm produced from the seeded samples of
    tests/golden/film_cases.py (make_golden_film.py), bit for bit,
  * the reference itself, live, where oracle/_ref exists (build container),
  * properties: a constant radiance field reconstructs to that constant, the box filter is the per-pixel mean."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
import film_cases as F


@pytest.mark.parametrize("case", range(len(F.CASES)))
def test_film_port_matches_pbrt_golden_bit_for_bit(case):
    name, kw, xres, yres, spp, window = F.CASES[case]
    flt = O.filter_args(name, **kw)
    cells, xy, rgb, px = F.make_samples(flt, xres, yres, spp, window, seed=100 + case)
    img = O.film(flt, cells, spp, xy, rgb, px)
    ref = load_golden("film")["film_%d" % case]
    assert img.shape == ref.shape
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref only exists in the build container")
def test_film_port_matches_pbrt_live():
    for n, name in enumerate(O.FILTERS):
        flt = O.filter_args(name)
        cells, xy, rgb, px = F.make_samples(flt, 31, 17, 2, (3, 1, 29, 15), seed=7 + n)
        img = O.film(flt, cells, 2, xy, rgb, px)
        ref = O.ref_film(flt, 31, 17, xy, rgb, px)
        assert np.array_equal(img.view(np.uint32), ref.view(np.uint32)), name


@pytest.mark.parametrize("name", list(O.FILTERS))
def test_constant_radiance_reconstructs_to_the_constant(name):
    flt = O.filter_args(name)
    cells, xy, rgb, px = F.make_samples(flt, 16, 12, 4, None, seed=3)
    rgb[:] = (0.25, 0.5, 2.0)
    img = O.film(flt, cells, 4, xy, rgb, px)
    assert np.abs(img - np.float32([0.25, 0.5, 2.0])).max() <= 2e-5


def test_box_filter_is_the_pixel_mean_and_halo_cells_do_not_reach_the_window():
    flt = O.filter_args("box")
    cells, xy, rgb, px = F.make_samples(flt, 16, 12, 4, None, seed=4)
    assert cells == (-1, -1, 17, 13)
    img = O.film(flt, cells, 4, xy, rgb, px)
    full = rgb.reshape(14, 18, 4, 3)[1:-1, 1:-1].astype(np.float64).mean(2)
    assert np.abs(img - full).max() <= 1e-5 * np.abs(full).max()


def test_filter_table_values():
    t = O.filter_table(O.filter_args("triangle")).reshape(16, 16)
    fx = (np.arange(16, dtype=np.float32) + np.float32(0.5)) * np.float32(2.0) / np.float32(16)
    assert np.array_equal(t, np.outer(np.float32(2) - fx, np.float32(2) - fx).astype(np.float32))
    m = O.filter_table(O.filter_args("mitchell")).reshape(16, 16)
    assert m[0, 0] > 0 and m[15, 15] > 0 and m[0, 12] < 0            # negative lobe of the Mitchell filter


LATTICE = [("box", {}), ("box", {"xwidth": 1.0, "ywidth": 0.75}), ("triangle", {}), ("mitchell", {"xwidth": 1.5, "ywidth": 2.25})]


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref only exists in the build container")
@pytest.mark.parametrize("name,kw", LATTICE)
def test_film_port_matches_pbrt_live_on_border_samples(name, kw):
    # samples exactly on cell borders, pixel centres and quarter positions: the integer edge of pbrt's extent rounding
    flt = O.filter_args(name, **kw)
    cells, xy, rgb, px = F.make_lattice_samples(flt, 13, 9, (1, 1, 12, 8))
    img = O.film(flt, cells, 16, xy, rgb, px)
    ref = O.ref_film(flt, 13, 9, xy, rgb, px)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32))


# ---- image-plane sample patterns -----------------------------------------------------------------------------------------
# known answers of the (0,2)-sequence with zero scrambles: radical inverse base 2, and the Sobol' second dimension in
# natural (not Gray-code) index order - direction numbers .1, .11, .101: n = 3 -> .1 ^ .11 = .01, n = 5 -> .101 ^ .1 = .001 ...
def test_ld_point_known_answers():
    assert [float(O.ld_point(n, 0, 0)[0]) for n in range(8)] == [0.0, 0.5, 0.25, 0.75, 0.125, 0.625, 0.375, 0.875]
    assert [float(O.ld_point(n, 0, 0)[1]) for n in range(8)] == [0.0, 0.5, 0.75, 0.25, 0.625, 0.125, 0.375, 0.875]


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref only exists in the build container")
def test_ld_point_and_strata_match_pbrt_live():
    rng = np.random.default_rng(2)
    for i in range(3000):
        n = i if i < 1024 else int(rng.integers(0, 1 << 24))
        s0, s1 = (int(v) for v in rng.integers(0, 1 << 32, 2, dtype=np.uint64))
        assert O.ld_point(n, s0, s1) == O.ref_ld_point(n, s0, s1)
    for nx, ny in ((1, 1), (2, 2), (4, 3), (3, 5)):
        # our stratified pattern with the jitter replaced by 0.5 = pbrt's StratifiedSample2D(jitter = false)
        c = O.ref_stratified_centres(nx, ny)
        dx, dy = np.float32(1) / np.float32(nx), np.float32(1) / np.float32(ny)
        k = np.arange(nx * ny)
        mine = np.stack([((k % nx).astype(np.float32) + np.float32(0.5)) * dx, ((k // nx).astype(np.float32) + np.float32(0.5)) * dy], -1)
        assert np.array_equal(mine.astype(np.float32), c)


@pytest.mark.parametrize("sampler,nx,ny,spp", [("random", 1, 1, 5), ("stratified", 3, 2, 6), ("lowdiscrepancy", 1, 1, 8)])
def test_sample_positions_structure(sampler, nx, ny, spp):
    win = (-2, -1, 14, 9)
    xy = O.sample_positions(sampler, 99, 12, 8, win, spp, nx, ny).reshape(10, 16, spp, 2)
    cx, cy = np.meshgrid(np.arange(win[0], win[2]), np.arange(win[1], win[3]))
    off = xy - np.stack([cx, cy], -1)[:, :, None, :].astype(np.float32)
    assert (off >= 0).all() and (off < 1).all()                     # every sample lies in its own cell
    assert len(np.unique(off.reshape(-1, 2), axis=0)) > 0.99 * off.size / 2       # halo and image cells get their own streams
    if sampler == "stratified":
        k = np.arange(spp)
        assert (np.floor(off[..., 0] * nx) == (k % nx)).all() and (np.floor(off[..., 1] * ny) == (k // nx)).all()
    if sampler == "lowdiscrepancy":
        # (0,2)-sequence: the first 8 points put exactly one point in every 1/8 column, 1/8 row and 1/4 x 1/2 box
        assert (np.sort(np.floor(off[..., 0] * 8), -1) == np.arange(8)).all()
        assert (np.sort(np.floor(off[..., 1] * 8), -1) == np.arange(8)).all()
        box = np.floor(off[..., 0] * 4) * 2 + np.floor(off[..., 1] * 2)
        assert (np.sort(box, -1) == np.arange(8)).all()
    # the same seed gives the same positions whatever the window
    part = O.sample_positions(sampler, 99, 12, 8, (3, 2, 7, 5), spp, nx, ny).reshape(3, 4, spp, 2)
    assert np.array_equal(part, xy[3:6, 5:9])
