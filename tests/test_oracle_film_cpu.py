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
