"""Shared by make_golden_film.py (build container, reference present) and the film tests: seeded sample sets for the
film reconstruction.  A case = (filter name, filter kwargs, xres, yres, spp, pixel window or None)."""
import numpy as np

CASES = [
    ("box", {}, 24, 16, 3, None),
    ("triangle", {}, 24, 16, 3, None),
    ("gaussian", {}, 24, 16, 3, None),
    ("mitchell", {}, 24, 16, 3, None),
    ("sinc", {}, 24, 16, 2, None),
    ("mitchell", {}, 24, 16, 3, (4, 2, 20, 14)),                 # crop window = an image band of a multi-GPU frame
    ("gaussian", {"xwidth": 1.5, "ywidth": 3.0, "alpha": 1.0}, 20, 12, 4, None),
    ("mitchell", {"B": 0.0, "C": 0.5, "xwidth": 3.0, "ywidth": 1.0}, 20, 12, 1, (0, 3, 20, 9)),
]


def halo(width):
    """sample cells beyond the pixel window that can still reach it: floor(width + 0.5)"""
    return int(np.floor(np.float32(width) + np.float32(0.5)))


def make_samples(flt, xres, yres, spp, window, seed):
    """cells (cx0, cy0, cx1, cy1), xy (n, 2), rgb (n, 3) in cell row-major order, spp jittered samples per cell"""
    x0, y0, x1, y1 = window if window is not None else (0, 0, xres, yres)
    hx, hy = halo(flt[1]), halo(flt[2])
    cells = (x0 - hx, y0 - hy, x1 + hx, y1 + hy)
    rng = np.random.default_rng(seed)
    cx, cy = np.meshgrid(np.arange(cells[0], cells[2]), np.arange(cells[1], cells[3]))
    base = np.stack([cx, cy], -1).reshape(-1, 1, 2).repeat(spp, 1).reshape(-1, 2).astype(np.float32)
    xy = base + rng.random(base.shape, dtype=np.float32)
    rgb = (rng.random((len(xy), 3), dtype=np.float32) * np.float32(3)) ** 2
    return cells, xy, rgb, (x0, y0, x1, y1)


def make_lattice_samples(flt, xres, yres, window=None):
    """Edge cases of the extent arithmetic: 16 samples per cell on the lattice {0, 0.25, 0.5, 0.75}^2 - positions on
    cell borders and pixel centres, where Ceil2Int / Floor2Int of imageX -+ width land exactly on integers."""
    x0, y0, x1, y1 = window if window is not None else (0, 0, xres, yres)
    hx, hy = halo(flt[1]), halo(flt[2])
    cells = (x0 - hx, y0 - hy, x1 + hx, y1 + hy)
    cx, cy = np.meshgrid(np.arange(cells[0], cells[2]), np.arange(cells[1], cells[3]))
    base = np.stack([cx, cy], -1).reshape(-1, 1, 2).astype(np.float32)
    off = np.stack(np.meshgrid(np.arange(4), np.arange(4)), -1).reshape(1, 16, 2).astype(np.float32) * np.float32(0.25)
    xy = (base + off).reshape(-1, 2)
    rng = np.random.default_rng(77)
    rgb = rng.random((len(xy), 3), dtype=np.float32)
    return cells, xy, rgb, (x0, y0, x1, y1)
