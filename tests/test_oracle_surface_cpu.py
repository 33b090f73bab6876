"""CPU tests: the C restatement of the hit attributes (oracle_photon.c tri_attr / port_surface_at) against the committed
golden that the reference's own pbrt-v2 code produced (tests/golden/make_golden_surface.py): per-vertex normals, uv,
reversed orientation.  Bar: every field bit-identical wherever both found the same primitive (exact-t ties excepted; BSDF
frame vectors equal up to the sign of a zero component, see compare())."""
import numpy as np
import pytest

from helpers import load_golden
from oracle import oracle as O
import make_golden_surface as G

FIELDS = ("p", "t", "ng", "u", "n", "v", "sn", "ray_eps", "tn")


def compare(got, ref):
    """count of differing words per field over the rays where both sides hit the same primitive; ties returned apart"""
    assert np.array_equal(got["prim"] >= 0, ref["prim"] >= 0)
    same = got["prim"] == ref["prim"]
    ties = int((~same).sum())
    bad = {f: int((np.ascontiguousarray(got[f][same]).view(np.uint32) != np.ascontiguousarray(ref[f][same]).view(np.uint32)).sum())
           for f in FIELDS}
    # the golden reads pbrt's private BSDF frame back through BSDF::LocalToWorld(unit axis) = axis * 1 + other * 0 + other * 0,
    # which turns a -0 component into +0: the frame vectors are compared as values (equal up to the sign of zero)
    for f in ("n", "sn", "tn"):
        bad[f] = int((got[f][same] != ref[f][same]).sum())
    return bad, ties


@pytest.mark.parametrize("case", G.CASES, ids=[c[0] for c in G.CASES])
def test_port_surface_matches_pbrt_golden(case):
    gold = load_golden("surface")[case[0]]
    ps = O.PortFullScene()
    G.fill(ps, case)
    got = ps.surface_at(G.rays())
    bad, ties = compare(got, gold)
    assert ties <= 4 and not any(bad.values()), (bad, ties)
    hit = gold["prim"] >= 0
    assert hit.sum() > 3000
    if case[1]:                                     # smooth normals: shading normal differs from the geometric one
        assert (np.abs(gold["n"][hit] - gold["ng"][hit]).max(axis=1) > 1e-3).mean() > 0.5
    if case[0] == "n_uv_rev":                       # reversed orientation flips both normals
        fwd = load_golden("surface")["n_uv"]
        assert np.array_equal(gold["ng"][hit], -fwd["ng"][hit]) and np.array_equal(gold["n"][hit], -fwd["n"][hit])


def test_golden_is_live_reference_output():
    """where oracle/_ref exists (build container) the committed golden must be what pbrt returns now"""
    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    rs = O.RefScene()
    for (P, idx, N, uv) in G.mesh():
        rs.add_mesh_attr(P, idx, N, uv, False)
    rs.build()
    live = rs.surface_at(G.rays(), 2)
    rs.close()
    assert live.tobytes() == load_golden("surface")["n_uv"].tobytes()
