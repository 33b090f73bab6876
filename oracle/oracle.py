"""TEST INFRASTRUCTURE ONLY: ctypes access to the two CPU oracles.

  * `port`  -- oracle/_build/liboracle.so, our plain-C restatement (oracle_port.c); always available.
  * `RefScene` -- oracle/_ref/libpbrt_oracle.so, the reference's own pbrt-v2 code compiled from
    /root/reference by oracle/Makefile (prebuilt files travel to the GPU box; sources do not).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(_HERE, "_build", "liboracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libpbrt_oracle.so")
REF_PBRT = os.path.join(_HERE, "_ref", "pbrt")


def build_port():
    subprocess.run(["make", "-C", _HERE, "port"], check=True, capture_output=True)


def build_ref():
    """Only possible where /root/reference exists (the build container)."""
    subprocess.run(["make", "-C", _HERE, "-j8", "ref"], check=True, capture_output=True)


def ref_available():
    return os.path.exists(REF_SO)


_port = None


def port():
    global _port
    if _port is None:
        if not os.path.exists(PORT_SO):
            build_port()
        lib = C.CDLL(PORT_SO)
        vp, i64 = C.c_void_p, C.c_int64
        lib.port_intersect_brute.argtypes = [C.c_int, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, C.c_int]
        lib.port_intersect_brute.restype = None
        lib.port_occluded_brute.argtypes = [C.c_int, vp, vp, vp, vp, i64, vp, vp, C.c_int]
        lib.port_occluded_brute.restype = None
        lib.port_lbvh_build.argtypes = [C.c_int] + [vp] * 8
        lib.port_prim_boxes.argtypes = [C.c_int, vp, vp, vp, vp]
        lib.port_prim_boxes.restype = None
        lib.port_sort_pairs.argtypes = [i64, vp, vp, C.c_int]
        lib.port_sort_pairs.restype = None
        _port = lib
    return _port


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class PortScene:
    """Flat scene in the same primitive numbering as b2rt (submission order)."""

    def __init__(self):
        self.verts = np.zeros((0, 3), np.float32)
        self.idx = np.zeros((0, 3), np.int32)
        self.kind = []
        self.analytic = []

    def add_mesh(self, P, idx):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        first = len(self.kind)
        t0 = len(self.idx)
        self.idx = np.concatenate([self.idx, idx + len(self.verts)]).astype(np.int32)
        self.verts = np.concatenate([self.verts, P]).astype(np.float32)
        self.kind.extend(range(t0, t0 + len(idx)))
        return first

    def _add_analytic(self, o2w, w2o, radius, height, inner, typ):
        rec = np.zeros(28, np.float32)
        rec[0:12] = np.asarray(o2w, np.float32).reshape(12)
        rec[12:24] = np.asarray(w2o, np.float32).reshape(12)
        rec[24:28] = (radius, height, inner, typ)
        self.kind.append(~len(self.analytic))
        self.analytic.append(rec)
        return len(self.kind) - 1

    def add_sphere(self, o2w, w2o, radius):
        return self._add_analytic(o2w, w2o, radius, 0.0, 0.0, 1)

    def add_disk(self, o2w, w2o, height, radius, inner=0.0):
        return self._add_analytic(o2w, w2o, height=height, radius=radius, inner=inner, typ=2)

    def _arrays(self):
        kind = np.asarray(self.kind, np.int32)
        an = np.concatenate(self.analytic).astype(np.float32) if self.analytic else np.zeros(28, np.float32)
        return kind, np.ascontiguousarray(self.verts), np.ascontiguousarray(self.idx), an

    def intersect(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        kind, verts, idx, an = self._arrays()
        t = np.zeros(n, np.float32)
        prim = np.zeros(n, np.int32)
        b1 = np.zeros(n, np.float32)
        b2 = np.zeros(n, np.float32)
        port().port_intersect_brute(len(kind), _p(kind), _p(verts), _p(idx), _p(an), n, _p(rays), _p(t), _p(prim),
                                    _p(b1), _p(b2), nthreads)
        return t, prim, b1, b2

    def occluded(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        kind, verts, idx, an = self._arrays()
        occ = np.zeros(n, np.uint8)
        port().port_occluded_brute(len(kind), _p(kind), _p(verts), _p(idx), _p(an), n, _p(rays), _p(occ), nthreads)
        return occ

    def tri_boxes(self):
        kind, verts, idx, _ = self._arrays()
        box = np.zeros((len(kind), 6), np.float32)
        port().port_prim_boxes(len(kind), _p(kind), _p(verts), _p(idx), _p(box))
        return box


def lbvh_build(box):
    """CPU restatement of the GPU builder on per-prim AABBs -> dict like Context.accel_download()."""
    box = np.ascontiguousarray(box, np.float32).reshape(-1, 6)
    n = len(box)
    out = {
        "morton": np.zeros(n, np.uint32), "prim_order": np.zeros(n, np.int32),
        "bin_parent": np.zeros(max(2 * n - 1, 0), np.int32), "bin_child": np.zeros(max(2 * (n - 1), 0), np.int32),
        "bin_box": np.zeros((max(2 * n - 1, 0), 6), np.float32), "collapse": np.zeros(max(n - 1, 1), np.uint8),
    }
    wide = np.zeros((max(n, 1), 32), np.uint32)
    nw = port().port_lbvh_build(n, _p(box), _p(out["morton"]), _p(out["prim_order"]), _p(out["bin_parent"]),
                                _p(out["bin_child"]), _p(out["bin_box"]), _p(out["collapse"]), _p(wide))
    out["wide_nodes"] = wide[:nw].copy()
    return out


def sort_pairs(keys, vals, key_bits=32):
    keys = np.ascontiguousarray(keys, np.uint32).copy()
    vals = np.ascontiguousarray(vals, np.uint32).copy()
    port().port_sort_pairs(len(keys), _p(keys), _p(vals), key_bits)
    return keys, vals


class RefScene:
    """The reference's own CPU path (pbrt-v2 BVHAccel + shapes) behind oracle/ref_capi.cpp."""

    def __init__(self):
        if not ref_available():
            raise FileNotFoundError(REF_SO + " (build with `make -C oracle ref` where /root/reference exists)")
        lib = C.CDLL(REF_SO)
        vp, i64, d = C.c_void_p, C.c_int64, C.c_double
        lib.ref_scene_create.restype = vp
        lib.ref_scene_add_mesh.argtypes = [vp, C.c_int, vp, C.c_int, vp]
        lib.ref_scene_add_sphere.argtypes = [vp, vp, vp, C.c_float]
        lib.ref_scene_add_disk.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_float]
        lib.ref_scene_build.argtypes = [vp]
        lib.ref_scene_build.restype = d
        lib.ref_scene_destroy.argtypes = [vp]
        lib.ref_scene_destroy.restype = None
        lib.ref_intersect.argtypes = [vp, i64, vp, vp, vp, vp, C.c_int]
        lib.ref_intersect.restype = d
        lib.ref_intersectp.argtypes = [vp, i64, vp, vp, C.c_int]
        lib.ref_intersectp.restype = d
        lib.ref_intersect_brute.argtypes = [vp, i64, vp, vp, vp, vp, C.c_int]
        lib.ref_intersect_brute.restype = d
        lib.ref_num_cores.restype = C.c_int
        self.lib = lib
        self.h = C.c_void_p(lib.ref_scene_create())
        self.build_seconds = None

    def add_mesh(self, P, idx):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        return self.lib.ref_scene_add_mesh(self.h, len(P), _p(P), len(idx), _p(idx))

    def add_sphere(self, o2w, w2o, radius):
        o2w = np.ascontiguousarray(o2w, np.float32).reshape(12)
        w2o = np.ascontiguousarray(w2o, np.float32).reshape(12)
        return self.lib.ref_scene_add_sphere(self.h, _p(o2w), _p(w2o), radius)

    def add_disk(self, o2w, w2o, height, radius, inner=0.0):
        o2w = np.ascontiguousarray(o2w, np.float32).reshape(12)
        w2o = np.ascontiguousarray(w2o, np.float32).reshape(12)
        return self.lib.ref_scene_add_disk(self.h, _p(o2w), _p(w2o), height, radius, inner)

    def build(self):
        self.build_seconds = self.lib.ref_scene_build(self.h)
        return self.build_seconds

    def intersect(self, rays, nthreads=1):
        """pbrt BVHAccel::Intersect.  Returns (t, prim, uv, seconds)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        t = np.zeros(n, np.float32)
        prim = np.zeros(n, np.int32)
        uv = np.zeros((n, 2), np.float32)
        sec = self.lib.ref_intersect(self.h, n, _p(rays), _p(t), _p(prim), _p(uv), nthreads)
        return t, prim, uv, sec

    def occluded(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = np.zeros(len(rays), np.uint8)
        sec = self.lib.ref_intersectp(self.h, len(rays), _p(rays), _p(occ), nthreads)
        return occ, sec

    def intersect_brute(self, rays, nthreads=1):
        """Exhaustive loop with pbrt's Shape::Intersect (min t, min id).  Returns (t, prim, nties)."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        t = np.zeros(n, np.float32)
        prim = np.zeros(n, np.int32)
        nties = C.c_int64()
        self.lib.ref_intersect_brute(self.h, n, _p(rays), _p(t), _p(prim), C.byref(nties), nthreads)
        return t, prim, nties.value

    def num_cores(self):
        return self.lib.ref_num_cores()

    def close(self):
        if self.h:
            self.lib.ref_scene_destroy(self.h)
            self.h = None

    __del__ = close
