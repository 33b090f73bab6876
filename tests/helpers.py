"""Shared by CPU and GPU tests: the golden cases (scene items regenerated from seeds)."""
import os
import numpy as np
from make_golden import case_scene, case_rays, fill  # noqa: F401  (tests/golden is on sys.path)

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ("terrain_small", "terrain_100k", "cornell")


def load_golden(name):
    return np.load(os.path.join(GOLDEN_DIR, name + ".npz"))


def fill_ctx(ctx, items):
    """Same items into a b2rt Context (argument order of Context.add_* differs from the oracle's)."""
    ctx.clear_scene()
    for it in items:
        if it[0] == "mesh":
            ctx.add_mesh(it[1], it[2])
        elif it[0] == "sphere":
            ctx.add_sphere(it[1], it[3], w2o=it[2])
        else:
            ctx.add_disk(it[1], it[3], it[4], it[5], w2o=it[2])


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)
