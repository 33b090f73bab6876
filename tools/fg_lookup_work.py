"""Work of the radiance-photon lookup (csrc/b2rt_fg.cuh fg_radiance_nearest) per class of query, counted on the HOST with
tests/hostemu (the same code the GPU runs, compiled for the CPU with counters): photons distance-tested and occupancy
cells looked at per query.  No GPU needed; used to choose the search order.    python tools/fg_lookup_work.py [n_paths]"""
import ctypes as C
import os
import subprocess
import sys
import time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O                     # noqa: E402
from cuda_raytrace_b200 import scenes as S         # noqa: E402

n_paths = int(sys.argv[1]) if len(sys.argv) > 1 else 300_000
so, src = os.path.join(ROOT, "build", "libfg_emu.so"), os.path.join(ROOT, "tests", "hostemu", "fg_emu.cu")
subprocess.run(["nvcc", "-O2", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-fmad=false",
                "-Wno-deprecated-gpu-targets", "-I" + os.path.join(ROOT, "include"), src, "-o", so], check=True)
emu = C.CDLL(so)
vp, f = C.c_void_p, C.c_float
emu.emu_radiance_nearest.argtypes = [vp, f, f, f, vp, vp, vp, vp, C.c_longlong, vp, vp, vp]
emu.emu_radiance_nearest.restype = None
sc = O.PortFullScene()
S.fill_scene(sc, S.cornell_scene(spheres=True, wide=1.78))
_, rad = sc.photon_trace_fg(3, 0, n_paths)
g = O.grid_build(rad)
lo, dims = np.asarray(g["origin"], np.float32), np.asarray(g["dims"], np.int32)
hi = lo + dims * np.float32(g["cell_size"])
slack = 1e-5 * float(max(np.abs(lo).max(), np.abs(hi).max())) + 1e-4 * g["cell_size"]
cs = np.ascontiguousarray(g["cell_start"], np.uint32)
cnt = np.diff(cs.astype(np.int64)).reshape(int(dims[2]), int(dims[1]), int(dims[0]))
cnt = np.pad(cnt, [(0, (-n) % 4) for n in cnt.shape])
z, y, x = cnt.shape
c1 = np.ascontiguousarray(cnt.reshape(z // 4, 4, y // 4, 4, x // 4, 4).sum((1, 3, 5)), np.uint32)
srt = np.ascontiguousarray(g["sorted"])
print("radiance photons %d, grid %s, cell %.4f" % (len(rad), list(dims), g["cell_size"]))
rng = np.random.default_rng(1)
nq = 2000


def run(name, p, n):
    q = np.zeros((len(p), 12), np.float32)
    q[:, 0:3], q[:, 3:6] = p, n
    idx, d2 = np.zeros(len(q), np.int32), np.zeros(len(q), np.float32)
    st = (C.c_ulonglong * 4)()
    emu.emu_stat(st, 1)
    t0 = time.perf_counter()
    emu.emu_radiance_nearest(lo.ctypes.data, g["cell_size"], g["inv_cell"], slack, dims.ctypes.data, srt.ctypes.data,
                             cs.ctypes.data, c1.ctypes.data, len(q), q.ctypes.data, idx.ctypes.data, d2.ctypes.data)
    t1 = time.perf_counter()
    emu.emu_stat(st, 1)
    pt = (C.c_ulonglong * 4)()
    emu.emu_phase_tests(pt, 1)
    print("      distance tests per query by phase: fine shells %.1f, sampling sweep %.1f, refinement %.1f, exact sweep %.1f"
          % tuple(v / len(q) for v in pt))
    _, ref = O.radiance_lookup(srt, q[:200])
    assert np.array_equal(idx[:200], ref)
    print("%-34s photons tested / query %9.1f   level-2 cells %7.1f   level-1 cells %7.1f   (sampled %6.1f)   found dist (cells) %5.1f   %.1f us/query (host)"
          % (name, st[0] / len(q), st[1] / len(q), st[2] / len(q), st[3] / len(q), float(np.sqrt(d2[idx >= 0]).mean() / g["cell_size"]),
             (t1 - t0) / len(q) * 1e6))


pick = rng.integers(0, len(rad), nq)
jit = rng.uniform(-1, 1, (nq, 3)).astype(np.float32) * np.float32(g["cell_size"])
run("on a lit matte surface, own normal", rad["p"][pick] + jit, O.oct_decode(rad["wi_oct"][pick]))
for name, c in (("glass sphere", (0.4, 0.35, 0.3)), ("mirror sphere", (-0.45, 0.35, -0.3))):
    v = rng.normal(size=(nq, 3))
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    run("on the " + name + ", outward normal", np.asarray(c) + 0.35 * v, v)
v = rng.normal(size=(nq, 3))
v /= np.linalg.norm(v, axis=1, keepdims=True)
run("inside the box, random normal", rng.uniform((-1.7, 0.05, -0.95), (1.7, 1.95, 0.95), (nq, 3)), v)

# ---- the lookups of real gather rays: cosine-sampled rays from lit surface points, exhaustive closest hit ------------------
ns = 6000
pick = rng.integers(0, len(rad), ns)
P0, N0 = rad["p"][pick].astype(np.float64), O.oct_decode(rad["wi_oct"][pick]).astype(np.float64)
u1, u2 = rng.random(ns), rng.random(ns)
rr, ph = np.sqrt(u1), 2 * np.pi * u2
t1 = np.where(np.abs(N0[:, :1]) > 0.5, np.array([[0.0, 1.0, 0.0]]), np.array([[1.0, 0.0, 0.0]]))
sn = np.cross(N0, t1); sn /= np.linalg.norm(sn, axis=1, keepdims=True)
tn = np.cross(N0, sn)
wi = sn * (rr * np.cos(ph))[:, None] + tn * (rr * np.sin(ph))[:, None] + N0 * np.sqrt(1 - u1)[:, None]
rays = np.zeros((ns, 8), np.float32)
rays[:, 0:3], rays[:, 3], rays[:, 4:7], rays[:, 7] = P0, 1e-3, wi, np.inf
t, prim, _, _ = sc.geo.intersect(rays)
hit = prim >= 0
kind = np.asarray(sc.geo.kind, np.int32)
hp = (P0 + wi * t[:, None].astype(np.float64))[hit]
hk = kind[prim[hit]]
ng = np.zeros_like(hp)
verts, idx = np.asarray(sc.geo.verts, np.float64).reshape(-1, 3), np.asarray(sc.geo.idx, np.int64).reshape(-1, 3)
tri = hk >= 0
v = verts[idx[hk[tri]]]
nt = np.cross(v[:, 2] - v[:, 0], v[:, 1] - v[:, 0])
ng[tri] = nt / np.linalg.norm(nt, axis=1, keepdims=True)
for c in ((0.4, 0.35, 0.3), (-0.45, 0.35, -0.3)):
    on = (~tri) & (np.abs(np.linalg.norm(hp - np.asarray(c), axis=1) - 0.35) < 1e-3)
    ng[on] = (hp[on] - np.asarray(c)) / 0.35
disk = (~tri) & (np.linalg.norm(ng, axis=1) == 0)
ng[disk] = (0, 0, 1)
flip = np.einsum("ij,ij->i", ng, -wi[hit]) < 0
ng[flip] *= -1
print("gather rays: %d of %d hit; %.1f %% of the hits on the spheres, %.1f %% on the light's disk"
      % (hit.sum(), ns, 100 * (~tri & ~disk).mean(), 100 * disk.mean()))
run("gather-ray hits on matte triangles", hp[tri], ng[tri])
if (~tri & ~disk).any():
    run("gather-ray hits on the spheres", hp[~tri & ~disk], ng[~tri & ~disk])
if disk.any():
    run("gather-ray hits on the light's disk", hp[disk], ng[disk])

if os.environ.get("FG_DEBUG"):
    sel = np.where(~tri & ~disk)[0][:60]
    for i in sel:
        q = np.zeros((1, 12), np.float32); q[0, 0:3], q[0, 3:6] = hp[i], ng[i]
        idx, d2 = np.zeros(1, np.int32), np.zeros(1, np.float32)
        st = (C.c_ulonglong * 4)(); emu.emu_stat(st, 1)
        emu.emu_radiance_nearest(lo.ctypes.data, g["cell_size"], g["inv_cell"], slack, dims.ctypes.data, srt.ctypes.data,
                                 cs.ctypes.data, c1.ctypes.data, 1, q.ctypes.data, idx.ctypes.data, d2.ctypes.data)
        emu.emu_stat(st, 1)
        ph = srt[idx[0]]
        print("p %s n %s -> photon p %s normal %s dist %.2f | tests %d l2 %d l1 %d sampled %d"
              % (np.round(hp[i], 2), np.round(ng[i], 2), np.round(ph["p"], 2), np.round(O.oct_decode(np.array([ph["wi_oct"]]))[0], 2),
                 np.sqrt(d2[0]), st[0], st[1], st[2], st[3]))
