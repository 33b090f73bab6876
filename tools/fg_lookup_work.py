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
    _, ref = O.radiance_lookup(srt, q[:200])
    assert np.array_equal(idx[:200], ref)
    print("%-34s photons tested / query %9.1f   level-2 cells %7.1f   level-1 cells %7.1f   found dist (cells) %5.1f   %.1f us/query (host)"
          % (name, st[0] / len(q), st[1] / len(q), st[2] / len(q), float(np.sqrt(d2[idx >= 0]).mean() / g["cell_size"]),
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
