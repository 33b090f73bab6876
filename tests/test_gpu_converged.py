"""Converged-image parity (-m gpu) against the reference's own CPU renderer on the BASELINE configs, at the north star's
bar: mean relative error <= 1 %, RMSE <= 0.5 % (exrdiff's RMS, pbrt-v2/src/tools/exrdiff.cpp:73-103).  The reference images
were rendered once in the build container by tools/converged_parity.py --ref (oracle/_ref/pbrt, the unmodified pbrt-v2:
16 independent realisations of 256 spp each, averaged) and are committed as tests/golden/converged.npz; the GPU side is
rendered here through the C ABI, 64 seeds averaged (0.2 s each; the error after 1, 2, 4 .. seeds is printed as `curve`).  Cfg1 exactly as BASELINE.json states it (4 spp, one 200 k photon map,
nused 50) is not a converged image on either side; its numbers are reported against pbrt's own noise floor, without a bar."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import converged_parity as CP   # noqa: E402

pytestmark = pytest.mark.gpu


def _have(case):
    return os.path.exists(CP.GOLD) and case in np.load(CP.GOLD)


@pytest.mark.parametrize("case", ["cfg1_converged", "cfg3_converged"])
def test_converged_image_within_north_star_bars(built, case):
    if not _have(case):
        pytest.skip("tests/golden/converged.npz has no %s (tools/converged_parity.py --ref)" % case)
    res, _ = CP.compare_gpu(case)
    m = res["metrics"]
    print(json.dumps(res))
    assert abs(m["exrdiff_avg_delta_pct"]) <= 0.5, m              # no bias
    assert m["mean_rel_err_pct"] <= 1.0, m                          # north star: 1 % mean relative error
    assert m["exrdiff_rms_pct"] <= 0.5, m                           # north star: 0.5 % RMSE (exrdiff definition)


def test_cfg1_as_stated_is_at_the_reference_noise_floor(built):
    """BASELINE configs[0] to the letter.  pbrt against pbrt with another pixel sampler (same photon map!) already differs
    by ~2.5 % mean / 7.8 % RMS; the GPU image, with its own photon map on top, must stay within 2.5x of that floor and be
    unbiased."""
    if not _have("cfg1_as_stated"):
        pytest.skip("tests/golden/converged.npz has no cfg1_as_stated")
    res, _ = CP.compare_gpu("cfg1_as_stated")
    m, floor = res["metrics"], res["pbrt_vs_pbrt_other_sampler"]
    print(json.dumps(res))
    assert abs(m["exrdiff_avg_delta_pct"]) <= 1.0, m
    assert m["mean_rel_err_pct"] <= 2.5 * floor["mean_rel_err_pct"], (m, floor)
    assert m["exrdiff_rms_pct"] <= 2.5 * floor["exrdiff_rms_pct"], (m, floor)
