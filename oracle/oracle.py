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
        self.vnormal = np.zeros((0, 3), np.float32)     # per-vertex attributes, zero where a mesh has none
        self.vuv = np.zeros((0, 2), np.float32)
        self.tri_flags = np.zeros(0, np.uint8)          # bit 0 normals, bit 1 uv, bit 2 reversed orientation

    def add_mesh(self, P, idx, N=None, uv=None, reverse=False):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        first = len(self.kind)
        t0 = len(self.idx)
        self.idx = np.concatenate([self.idx, idx + len(self.verts)]).astype(np.int32)
        self.verts = np.concatenate([self.verts, P]).astype(np.float32)
        self.kind.extend(range(t0, t0 + len(idx)))
        fl = (1 if N is not None else 0) | (2 if uv is not None else 0) | (4 if reverse else 0)
        self.vnormal = np.concatenate([self.vnormal, np.zeros((len(P), 3), np.float32) if N is None else
                                       np.asarray(N, np.float32).reshape(-1, 3)]).astype(np.float32)
        self.vuv = np.concatenate([self.vuv, np.zeros((len(P), 2), np.float32) if uv is None else
                                   np.asarray(uv, np.float32).reshape(-1, 2)]).astype(np.float32)
        self.tri_flags = np.concatenate([self.tri_flags, np.full(len(idx), fl, np.uint8)])
        return first

    def _attr_arrays(self):
        """(vnormal, vuv, tri_flags) or (None, None, None) when no mesh has attributes"""
        if not self.tri_flags.any():
            return None, None, None
        return np.ascontiguousarray(self.vnormal), np.ascontiguousarray(self.vuv), np.ascontiguousarray(self.tri_flags)

    def _add_analytic(self, o2w, w2o, radius, height, inner, typ, phimax=0.0):
        rec = np.zeros(32, np.float32)                 # PORT_ANALYTIC_STRIDE
        rec[0:12] = np.asarray(o2w, np.float32).reshape(12)
        rec[12:24] = np.asarray(w2o, np.float32).reshape(12)
        rec[24:28] = (radius, height, inner, typ)
        rec[28] = phimax                                # radians, 0 = full turn
        self.kind.append(~len(self.analytic))
        self.analytic.append(rec)
        return len(self.kind) - 1

    @staticmethod
    def _phimax_field(phimax_deg):
        """degrees -> the record's value: Radians(Clamp(pm, 0, 360)) in binary32 as pbrt forms it, 0 for the full turn"""
        if phimax_deg is None:
            return 0.0
        pm = np.float32(min(max(phimax_deg, 0.0), 360.0))
        return 0.0 if pm >= 360.0 else float((np.float32(np.pi) / np.float32(180.0)) * pm)

    def add_sphere(self, o2w, w2o, radius, zmin=None, zmax=None, phimax=None):
        """zmin / zmax / phimax: pbrt's partial sphere; clamped as Sphere::Sphere does, (0, 0) in the record = no z clipping"""
        z0, z1 = 0.0, 0.0
        if zmin is not None or zmax is not None:
            a, b = (-radius if zmin is None else zmin), (radius if zmax is None else zmax)
            z0 = float(np.clip(np.float32(min(a, b)), -np.float32(radius), np.float32(radius)))
            z1 = float(np.clip(np.float32(max(a, b)), -np.float32(radius), np.float32(radius)))
            if z0 <= -radius and z1 >= radius:
                z0 = z1 = 0.0
        return self._add_analytic(o2w, w2o, radius, z0, z1, 1, self._phimax_field(phimax))

    def add_disk(self, o2w, w2o, height, radius, inner=0.0, phimax=None):
        return self._add_analytic(o2w, w2o, height=height, radius=radius, inner=inner, typ=2, phimax=self._phimax_field(phimax))

    def _arrays(self):
        kind = np.asarray(self.kind, np.int32)
        an = np.concatenate(self.analytic).astype(np.float32) if self.analytic else np.zeros(32, np.float32)
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


def xf_point32(m, x, y, z):
    """pbrt Transform::operator()(Point) for an affine 3x4 matrix in binary32, one rounding per operation, summed left to
    right (core/transform.h:  m00*x + m01*y + m02*z + m03) -- the same expression as csrc/b2rt_device.cuh xf_point"""
    m = np.asarray(m, np.float32).reshape(12)
    return (m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7], m[8] * x + m[9] * y + m[10] * z + m[11])


def xf_vec32(m, x, y, z):
    m = np.asarray(m, np.float32).reshape(12)
    return (m[0] * x + m[1] * y + m[2] * z, m[4] * x + m[5] * y + m[6] * z, m[8] * x + m[9] * y + m[10] * z)


class PortInstancedScene:
    """Oracle of the two-level scene (b2rt_object_begin / b2rt_add_instance), layered over the exhaustive C loop: pbrt's
    TransformedPrimitive::Intersect (core/primitive.cpp:95-103) takes the ray into the instance's space with
    Transform::operator()(Ray) -- origin as a point, direction as a vector, tmin / tmax unchanged -- and intersects the
    object there.  Ids as in include/b2rt.h: items (world primitives and instances, submission order) first, then the
    primitives of instance 0, instance 1, ...  Winner = (min t, then min id)."""

    def __init__(self):
        self.world = PortScene()
        self.objects = []
        self.instances = []          # (object, o2w, w2o, item id)
        self._cur = None
        self.n_items = 0

    def object_begin(self):
        self.objects.append(PortScene())
        self._cur = self.objects[-1]
        return len(self.objects) - 1

    def object_end(self):
        self._cur = None

    def _target(self):
        return self._cur if self._cur is not None else self.world

    def add_mesh(self, P, idx, **kw):
        if self._cur is not None:
            return self._cur.add_mesh(P, idx, **kw)
        first = self.n_items
        self.world.add_mesh(P, idx, **kw)
        self._item_of_world = getattr(self, "_item_of_world", [])
        self._item_of_world.extend(range(first, first + len(np.asarray(idx).reshape(-1, 3))))
        self.n_items += len(np.asarray(idx).reshape(-1, 3))
        return first

    def _add_one(self, fn, *a):
        if self._cur is not None:
            return fn(self._cur, *a)
        fn(self.world, *a)
        self._item_of_world = getattr(self, "_item_of_world", [])
        self._item_of_world.append(self.n_items)
        self.n_items += 1
        return self.n_items - 1

    def add_sphere(self, o2w, w2o, radius):
        return self._add_one(PortScene.add_sphere, o2w, w2o, radius)

    def add_disk(self, o2w, w2o, height, radius, inner=0.0):
        return self._add_one(PortScene.add_disk, o2w, w2o, height, radius, inner)

    def add_instance(self, obj, o2w, w2o):
        self.instances.append((obj, np.asarray(o2w, np.float32).reshape(12), np.asarray(w2o, np.float32).reshape(12), self.n_items))
        self.n_items += 1
        return len(self.instances) - 1

    def instance_bases(self):
        base, out = self.n_items, []
        for obj, _, _, _ in self.instances:
            out.append(base)
            base += len(self.objects[obj].kind)
        return out

    def instance_rays(self, k, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        _, _, w2o, _ = self.instances[k]
        r = rays.copy()
        r[:, 0], r[:, 1], r[:, 2] = xf_point32(w2o, rays[:, 0], rays[:, 1], rays[:, 2])
        r[:, 4], r[:, 5], r[:, 6] = xf_vec32(w2o, rays[:, 4], rays[:, 5], rays[:, 6])
        return r

    def intersect(self, rays, nthreads=1):
        """(t, global prim id, b1, b2, instance index or -1)"""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        t = np.full(n, np.inf, np.float32); prim = np.full(n, -1, np.int32)
        b1 = np.zeros(n, np.float32); b2 = np.zeros(n, np.float32); inst = np.full(n, -1, np.int32)
        if self.world.kind:
            wt, wp, wb1, wb2 = self.world.intersect(rays, nthreads)
            items = np.asarray(self._item_of_world, np.int32)
            hit = wp >= 0
            t[hit], prim[hit], b1[hit], b2[hit] = wt[hit], items[wp[hit]], wb1[hit], wb2[hit]
        for k, base in enumerate(self.instance_bases()):
            it, ip, ib1, ib2 = self.objects[self.instances[k][0]].intersect(self.instance_rays(k, rays), nthreads)
            gp = ip + base
            better = (ip >= 0) & ((it < t) | ((it == t) & ((prim < 0) | (gp < prim))))
            t[better], prim[better], b1[better], b2[better], inst[better] = it[better], gp[better], ib1[better], ib2[better], k
        return t, prim, b1, b2, inst

    def occluded(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        occ = self.world.occluded(rays, nthreads) if self.world.kind else np.zeros(len(rays), np.uint8)
        for k in range(len(self.instances)):
            occ |= self.objects[self.instances[k][0]].occluded(self.instance_rays(k, rays), nthreads)
        return occ


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
        lib.ref_scene_add_mesh_attr.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp, vp, C.c_int]
        lib.ref_surface.argtypes = [vp, i64, vp, vp, C.c_int]
        lib.ref_surface.restype = None
        self.lib = lib
        self.h = C.c_void_p(lib.ref_scene_create())
        self.build_seconds = None

    def add_mesh(self, P, idx):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        return self.lib.ref_scene_add_mesh(self.h, len(P), _p(P), len(idx), _p(idx))

    def add_mesh_attr(self, P, idx, N=None, uv=None, reverse=False):
        """pbrt TriangleMesh with "N" / "uv" (identity object-to-world, so N is in world space) and ReverseOrientation"""
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        N = None if N is None else np.ascontiguousarray(N, np.float32).reshape(-1, 3)
        uv = None if uv is None else np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
        return self.lib.ref_scene_add_mesh_attr(self.h, len(P), _p(P), len(idx), _p(idx), _p(N), _p(uv), 1 if reverse else 0)

    def surface_at(self, rays, nthreads=1):
        """pbrt's own hit attributes: BVHAccel::Intersect, Intersection::GetBSDF (Triangle::GetShadingGeometry) and the
        BSDF's (nn, sn, tn) frame.  SURFACE_DTYPE per ray."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays), SURFACE_DTYPE)
        self.lib.ref_surface(self.h, len(rays), _p(rays), _p(out), nthreads)
        return out

    def add_sphere(self, o2w, w2o, radius, zmin=None, zmax=None, phimax=None):
        o2w = np.ascontiguousarray(o2w, np.float32).reshape(12)
        w2o = np.ascontiguousarray(w2o, np.float32).reshape(12)
        if zmin is not None or zmax is not None or phimax is not None:
            self.lib.ref_scene_add_sphere_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float]
            return self.lib.ref_scene_add_sphere_ex(self.h, _p(o2w), _p(w2o), radius, -radius if zmin is None else zmin,
                                                    radius if zmax is None else zmax, 360.0 if phimax is None else phimax)
        return self.lib.ref_scene_add_sphere(self.h, _p(o2w), _p(w2o), radius)

    def add_disk(self, o2w, w2o, height, radius, inner=0.0, phimax=None):
        o2w = np.ascontiguousarray(o2w, np.float32).reshape(12)
        w2o = np.ascontiguousarray(w2o, np.float32).reshape(12)
        if phimax is not None:
            self.lib.ref_scene_add_disk_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float]
            return self.lib.ref_scene_add_disk_ex(self.h, _p(o2w), _p(w2o), height, radius, inner, phimax)
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

    # ---- object instances: pbrt's own ObjectInstance path (TransformedPrimitive over the object's BVHAccel) --------
    def object_begin(self):
        return self.lib.ref_scene_object_begin(self.h)

    def object_end(self):
        self.lib.ref_scene_object_end.restype = None
        self.lib.ref_scene_object_end(self.h)

    def add_instance(self, obj, o2w, w2o):
        o2w = np.ascontiguousarray(o2w, np.float32).reshape(12)
        w2o = np.ascontiguousarray(w2o, np.float32).reshape(12)
        self.lib.ref_scene_add_instance.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        return self.lib.ref_scene_add_instance(self.h, obj, _p(o2w), _p(w2o))

    def intersect_inst(self, rays, nthreads=1):
        """(t, prim, inst): prim is local to the object where inst >= 0"""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        n = len(rays)
        t, prim, inst = np.zeros(n, np.float32), np.zeros(n, np.int32), np.zeros(n, np.int32)
        self.lib.ref_intersect_inst.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        self.lib.ref_intersect_inst.restype = C.c_double
        self.lib.ref_intersect_inst(self.h, n, _p(rays), _p(t), _p(prim), _p(inst), nthreads)
        return t, prim, inst

    def surface_at_inst(self, rays, nthreads=1):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays), SURFACE_DTYPE)
        inst = np.zeros(len(rays), np.int32)
        self.lib.ref_surface_inst.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        self.lib.ref_surface_inst.restype = None
        self.lib.ref_surface_inst(self.h, len(rays), _p(rays), _p(out), _p(inst), nthreads)
        return out, inst

    def close(self):
        if self.h:
            self.lib.ref_scene_destroy(self.h)
            self.h = None

    __del__ = close


# ---- photon pass / grid / estimator restatements (oracle_photon.c) ----------------------------------------------------
PHOTON_DTYPE = np.dtype([("p", np.float32, 3), ("wi_oct", np.uint32), ("alpha", np.float32, 3), ("path", np.uint32)])
# == b2rt_surface (include/b2rt.h): hit point, t, geometric normal, u, shading normal (BSDF nn), v, sn, ray epsilon, tn, prim
SURFACE_DTYPE = np.dtype([("p", np.float32, 3), ("t", np.float32), ("ng", np.float32, 3), ("u", np.float32),
                          ("n", np.float32, 3), ("v", np.float32), ("sn", np.float32, 3), ("ray_eps", np.float32),
                          ("tn", np.float32, 3), ("prim", np.int32)])


class _CScene(C.Structure):
    _fields_ = [("nprims", C.c_int32), ("kind", C.c_void_p), ("verts", C.c_void_p), ("idx", C.c_void_p),
                ("analytic", C.c_void_p), ("prim_material", C.c_void_p), ("nmaterials", C.c_int32),
                ("materials", C.c_void_p), ("nlights", C.c_int32), ("lights", C.c_void_p),
                ("vnormal", C.c_void_p), ("vuv", C.c_void_p), ("tri_flags", C.c_void_p),
                ("prim_light", C.c_void_p), ("light_nsamples", C.c_void_p),
                ("light_ext", C.c_void_p), ("world_bounds", C.c_void_p)]


def _bind_photon():
    lib = port()
    if getattr(lib, "_photon_bound", False):
        return lib
    vp, i64, f = C.c_void_p, C.c_int64, C.c_float
    lib.port_photon_trace.argtypes = [C.POINTER(_CScene), C.c_uint64, i64, i64, C.c_int, C.c_int, vp, i64,
                                      C.POINTER(i64), vp, i64, C.POINTER(i64)]
    lib.port_grid_build.argtypes = [i64, vp, f, f, vp, vp, C.POINTER(f), C.POINTER(f), vp, vp, vp]
    lib.port_grid_build.restype = None
    lib.port_gather.argtypes = [i64, vp, i64, i64, vp, C.c_int, f, vp, vp, vp, vp]
    lib.port_gather.restype = None
    lib.port_photon_trace_fg.argtypes = [C.POINTER(_CScene), C.c_uint64, i64, i64, C.c_int, vp, i64, C.POINTER(i64), vp, i64,
                                         C.POINTER(i64)]
    lib.port_irradiance.argtypes = [i64, vp, i64, i64, vp, C.c_int, f, vp]
    lib.port_irradiance.restype = None
    lib.port_radiance_lookup.argtypes = [i64, vp, i64, vp, vp, vp]
    lib.port_radiance_lookup.restype = None
    lib.port_final_gather.argtypes = [C.POINTER(_CScene), i64, vp, i64, vp, i64, vp, C.c_int, f, C.c_uint64, vp]
    lib.port_final_gather.restype = None
    lib.port_camera_samples.argtypes = [C.POINTER(_CScene), C.c_uint64, vp, i64, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                        C.c_int, vp, vp, vp]
    lib.port_camera_samples.restype = None
    lib.port_surface_at.argtypes = [C.POINTER(_CScene), i64, vp, vp]
    lib.port_surface_at.restype = None
    lib.port_sincos_quarter.argtypes = [f, C.POINTER(f), C.POINTER(f)]
    lib.port_sincos_quarter.restype = None
    lib.port_philox.argtypes = [C.c_uint32] * 6 + [vp]
    lib.port_philox.restype = None
    lib._photon_bound = True
    return lib


class PortFullScene:
    """Scene with materials and lights, same method signatures as cuda_raytrace_b200.b2rt.Context so that
    scenes.fill_scene() fills either."""

    def __init__(self):
        self.geo = PortScene()
        self.prim_material = []
        self.prim_light = []
        self.materials = []
        self.lights = []
        self.light_nsamples = []
        self.light_ext = []
        self.world_bounds = None      # scene box for distant lights: set from the GPU's accel_info().bounds (set_world_bounds)

    def add_material(self, mtype, rgb, kt=None, ior=1.5):
        kt = (1.0, 1.0, 1.0) if kt is None else kt
        self.materials.append([float(mtype), *map(float, rgb), *map(float, kt), float(ior)])
        return len(self.materials) - 1

    def add_mesh(self, P, idx, material=-1, arealight=-1, N=None, uv=None, reverse=False):
        first = self.geo.add_mesh(P, idx, N, uv, reverse)
        self.prim_material.extend([material] * len(np.asarray(idx).reshape(-1, 3)))
        self.prim_light.extend([arealight] * len(np.asarray(idx).reshape(-1, 3)))
        return first

    def add_sphere(self, o2w, radius, w2o=None, material=-1, arealight=-1, zmin=None, zmax=None, phimax=None):
        self.prim_material.append(material)
        self.prim_light.append(arealight)
        return self.geo.add_sphere(o2w, w2o, radius, zmin, zmax, phimax)

    def add_disk(self, o2w, height, radius, inner=0.0, w2o=None, material=-1, arealight=-1, phimax=None):
        self.prim_material.append(material)
        self.prim_light.append(arealight)
        return self.geo.add_disk(o2w, w2o, height, radius, inner, phimax)

    def add_light(self, ltype, rgb, p=None, prim=-1, nsamples=1):
        p = (0.0, 0.0, 0.0) if p is None else p
        self.lights.append([float(ltype), *map(float, p), *map(float, rgb), float(prim)])
        self.light_nsamples.append(int(nsamples))
        self.light_ext.append(np.zeros(24, np.float32))
        return len(self.lights) - 1

    def add_spot_light(self, l2w, w2l, rgb, cos_total, cos_falloff):
        l2w, w2l = np.asarray(l2w, np.float32).reshape(3, 4), np.asarray(w2l, np.float32).reshape(3, 4)
        ext = np.zeros(24, np.float32)
        ext[0:9], ext[9:18], ext[18], ext[19] = l2w[:, :3].reshape(-1), w2l[:, :3].reshape(-1), cos_total, cos_falloff
        self.lights.append([2.0, *map(float, l2w[:, 3]), *map(float, rgb), -1.0])
        self.light_nsamples.append(1)
        self.light_ext.append(ext)
        return len(self.lights) - 1

    def add_distant_light(self, dir_to_light, rgb):
        ext = np.zeros(24, np.float32)
        ext[20:23] = np.asarray(dir_to_light, np.float32)
        self.lights.append([3.0, 0.0, 0.0, 0.0, *map(float, rgb), -1.0])
        self.light_nsamples.append(1)
        self.light_ext.append(ext)
        return len(self.lights) - 1

    def set_world_bounds(self, bounds6):
        self.world_bounds = np.ascontiguousarray(bounds6, np.float32).reshape(6)

    def _cscene(self):
        kind, verts, idx, an = self.geo._arrays()
        pm = np.asarray(self.prim_material, np.int32)
        mats = np.asarray(self.materials, np.float32).reshape(-1, 8)
        lights = np.asarray(self.lights, np.float32).reshape(-1, 8)
        vn, vuv, tf = self.geo._attr_arrays()
        pl = np.asarray(self.prim_light, np.int32)
        lns = np.asarray(self.light_nsamples, np.int32)
        lext = np.ascontiguousarray(np.concatenate(self.light_ext)) if self.light_ext else None
        wb = self.world_bounds
        keep = (kind, verts, idx, an, pm, mats, lights, vn, vuv, tf, pl, lns, lext, wb)
        cs = _CScene(len(kind), _p(kind).value, _p(verts).value, _p(idx).value, _p(an).value,
                     _p(pm).value if len(pm) else None, len(mats),
                     _p(mats).value if len(mats) else None, len(lights), _p(lights).value if len(lights) else None,
                     _p(vn).value if vn is not None else None, _p(vuv).value if vuv is not None else None,
                     _p(tf).value if tf is not None else None, _p(pl).value if len(pl) else None,
                     _p(lns).value if len(lns) else None, _p(lext).value if lext is not None else None,
                     _p(wb).value if wb is not None else None)
        return cs, keep

    def camera_samples(self, seed, cam12, xy, gpix, ks, sampler_type=0, max_specular_depth=5, direct=True, slots=0):
        """port_camera_samples: per-sample camera-pass radiance (n, 3), leaf counts (n,), leaf queries (n, slots, 12)"""
        lib = _bind_photon()
        cs, keep = self._cscene()
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        n = len(xy)
        gpix, ks = np.ascontiguousarray(gpix, np.uint32), np.ascontiguousarray(ks, np.uint32)
        cam12 = np.ascontiguousarray(cam12, np.float32).reshape(12)
        rgb, nq = np.zeros((n, 3), np.float32), np.zeros(n, np.uint32)
        q = np.zeros((n, max(slots, 1), 12), np.float32)
        lib.port_camera_samples(C.byref(cs), seed, _p(cam12), n, _p(xy), _p(gpix), _p(ks), sampler_type, max_specular_depth,
                                1 if direct else 0, 1 if slots > 0 else 0, max(slots, 1), _p(rgb), _p(nq), _p(q))
        del keep
        return rgb, nq, q

    def surface_at(self, rays):
        """port_surface_at: closest hit + hit attributes, SURFACE_DTYPE per ray"""
        lib = _bind_photon()
        cs, keep = self._cscene()
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays), SURFACE_DTYPE)
        lib.port_surface_at(C.byref(cs), len(rays), _p(rays), _p(out))
        del keep
        return out

    def photon_trace(self, seed, first_path, n_paths, max_depth=5, caustic=True, indirect=True):
        lib = _bind_photon()
        cs, keep = self._cscene()
        cap = n_paths * max_depth + 1
        oc, oi = np.zeros(cap, PHOTON_DTYPE), np.zeros(cap, PHOTON_DTYPE)
        nc, ni = C.c_int64(0), C.c_int64(0)
        rc = lib.port_photon_trace(C.byref(cs), seed, first_path, n_paths, max_depth,
                                   (1 if caustic else 0) | (2 if indirect else 0), _p(oc), cap, C.byref(nc), _p(oi), cap,
                                   C.byref(ni))
        assert rc == 0, rc
        del keep
        return oc[: nc.value].copy(), oi[: ni.value].copy()


def _pft_fg(self, seed, first_path, n_paths, max_depth=5, limits=(-1, -1)):
    """(direct photons, radiance photons) of the path range: port_photon_trace_fg.  limits = (caustic_paths, indirect_paths)
    of pbrt's depositedPhoton rule (-1: none)."""
    lib = _bind_photon()
    lib.port_fg_deposit_limits.argtypes = [C.c_int64, C.c_int64]
    lib.port_fg_deposit_limits.restype = None
    big = 2 ** 63 - 1
    lib.port_fg_deposit_limits(big if limits[0] < 0 else limits[0], big if limits[1] < 0 else limits[1])
    cs, keep = self._cscene()
    cap = n_paths * max_depth + 1
    od, orr = np.zeros(cap, PHOTON_DTYPE), np.zeros(cap, PHOTON_DTYPE)
    nd, nr = C.c_int64(0), C.c_int64(0)
    rc = lib.port_photon_trace_fg(C.byref(cs), seed, first_path, n_paths, max_depth, _p(od), cap, C.byref(nd), _p(orr), cap,
                                  C.byref(nr))
    lib.port_fg_deposit_limits(big, big)
    assert rc == 0, rc
    del keep
    return od[: nd.value].copy(), orr[: nr.value].copy()


def _pfinal_gather(self, ind_sorted, rad_sorted, queries, gather_samples, cos_angle, seed):
    """port_final_gather: (nq, 3) indirect radiance of the queries (12 floats each, pad = ray eps, pixel id, sample word)."""
    lib = _bind_photon()
    cs, keep = self._cscene()
    ind = np.ascontiguousarray(ind_sorted, PHOTON_DTYPE)
    rad = np.ascontiguousarray(rad_sorted, PHOTON_DTYPE)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    rgb = np.zeros((len(q), 3), np.float32)
    lib.port_final_gather(C.byref(cs), len(ind), _p(ind), len(rad), _p(rad), len(q), _p(q), max(1, gather_samples // 2),
                          float(cos_angle), seed, _p(rgb))
    del keep
    return rgb


PortFullScene.photon_trace_fg = _pft_fg
PortFullScene.final_gather = _pfinal_gather


def irradiance(sorted_photons, n_paths, queries, k, max_dist2):
    """port_irradiance (pbrt EPhoton): (nq, 3)."""
    lib = _bind_photon()
    sp = np.ascontiguousarray(sorted_photons, PHOTON_DTYPE)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    E = np.zeros((len(q), 3), np.float32)
    lib.port_irradiance(len(sp), _p(sp), n_paths, len(q), _p(q), k, max_dist2, _p(E))
    return E


def radiance_lookup(radiance_photons, queries):
    """port_radiance_lookup: (Lo (nq, 3), index or -1) of the nearest photon whose normal faces the query's."""
    lib = _bind_photon()
    rp = np.ascontiguousarray(radiance_photons, PHOTON_DTYPE)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    rgb, idx = np.zeros((len(q), 3), np.float32), np.zeros(len(q), np.int32)
    lib.port_radiance_lookup(len(rp), _p(rp), len(q), _p(q), _p(rgb), _p(idx))
    return rgb, idx


def grid_build(photons, cell_size=0.0, target_per_cell=0.0):
    lib = _bind_photon()
    photons = np.ascontiguousarray(photons, PHOTON_DTYPE)
    n = len(photons)
    origin, dims = np.zeros(3, np.float32), np.zeros(3, np.int32)
    cell, inv = C.c_float(), C.c_float()
    lib.port_grid_build(n, _p(photons), cell_size, target_per_cell, _p(origin), _p(dims), C.byref(cell), C.byref(inv),
                        None, None, None)
    ncell = int(dims[0]) * int(dims[1]) * int(dims[2])
    keys, order, cs = np.zeros(n, np.uint32), np.zeros(n, np.uint32), np.zeros(ncell + 1, np.uint32)
    lib.port_grid_build(n, _p(photons), cell_size, target_per_cell, _p(origin), _p(dims), C.byref(cell), C.byref(inv),
                        _p(keys), _p(order), _p(cs))
    return {"origin": origin, "dims": dims, "cell_size": cell.value, "inv_cell": inv.value, "keys": keys,
            "order": order, "cell_start": cs, "sorted": photons[order]}


def gather(sorted_photons, n_paths, queries, k, max_dist2, want_sets=False):
    """queries: (nq,12) float32 {p, n, kd, pad}.  Returns rgb, r2, found[, idx]."""
    lib = _bind_photon()
    sp = np.ascontiguousarray(sorted_photons, PHOTON_DTYPE)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    nq = len(q)
    rgb, r2, found = np.zeros((nq, 3), np.float32), np.zeros(nq, np.float32), np.zeros(nq, np.int32)
    idx = np.zeros((nq, max(k, 1)), np.int32) if want_sets else None
    lib.port_gather(len(sp), _p(sp), n_paths, nq, _p(q), k, max_dist2, _p(rgb), _p(r2), _p(found), _p(idx))
    return (rgb, r2, found, idx) if want_sets else (rgb, r2, found)


def sincos_quarter(q):
    lib = _bind_photon()
    s, c = C.c_float(), C.c_float()
    lib.port_sincos_quarter(float(q), C.byref(s), C.byref(c))
    return s.value, c.value


class RefPhotonScene:
    """The reference's own CPU photon mapper (pbrt-v2 PhotonIntegrator) behind oracle/ref_photon.cpp; method
    signatures follow b2rt.Context so scenes.fill_scene() fills it."""

    def __init__(self):
        if not ref_available():
            raise FileNotFoundError(REF_SO)
        lib = C.CDLL(REF_SO)
        vp, i64, f, i = C.c_void_p, C.c_int64, C.c_float, C.c_int
        lib.ref_pm_create.restype = vp
        lib.ref_pm_add_material.argtypes = [vp, i, vp, vp, f]
        lib.ref_pm_add_mesh.argtypes = [vp, i, vp, i, vp, i]
        lib.ref_pm_add_sphere.argtypes = [vp, vp, vp, f, i]
        lib.ref_pm_add_disk.argtypes = [vp, vp, vp, f, f, f, i, vp, i]
        lib.ref_pm_add_point_light.argtypes = [vp, vp, vp]
        lib.ref_pm_build.argtypes = [vp]
        lib.ref_pm_shoot.argtypes = [vp, i, i, i, i, vp, i64, vp, i64, vp, vp]
        lib.ref_pm_lphoton.argtypes = [i64, vp, i, i64, vp, i, f, vp]
        lib.ref_pm_lphoton.restype = None
        lib.ref_pm_ephoton.argtypes = [i64, vp, i, i64, vp, i, f, vp]
        lib.ref_pm_ephoton.restype = None
        lib.ref_pm_radiance_lookup.argtypes = [i64, vp, i64, vp, vp]
        lib.ref_pm_radiance_lookup.restype = None
        self.lib = lib
        self.h = C.c_void_p(lib.ref_pm_create())
        self.nprims = 0
        self.disks = {}          # first prim -> deferred disk args (a disk becomes an area light in add_light)
        self.built = False

    def add_material(self, mtype, rgb, kt=None, ior=1.5):
        a = np.asarray(rgb, np.float32)
        b = np.asarray((1, 1, 1) if kt is None else kt, np.float32)
        return self.lib.ref_pm_add_material(self.h, int(mtype), _p(a), _p(b), float(ior))

    def add_mesh(self, P, idx, material=-1, arealight=-1):
        P = np.ascontiguousarray(P, np.float32).reshape(-1, 3)
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 3)
        self.lib.ref_pm_add_mesh(self.h, len(P), _p(P), len(idx), _p(idx), material)
        first = self.nprims
        self.nprims += len(idx)
        return first

    def add_sphere(self, o2w, radius, w2o=None, material=-1, arealight=-1):
        a, b = np.ascontiguousarray(o2w, np.float32), np.ascontiguousarray(w2o, np.float32)
        self.lib.ref_pm_add_sphere(self.h, _p(a), _p(b), radius, material)
        self.nprims += 1
        return self.nprims - 1

    def add_disk(self, o2w, height, radius, inner=0.0, w2o=None, material=-1, arealight=-1):
        args = (np.ascontiguousarray(o2w, np.float32), np.ascontiguousarray(w2o, np.float32), height, radius, inner,
                material)
        if arealight >= 0:
            self.disks[self.nprims] = args      # created in add_light, when L is known
        else:
            self.lib.ref_pm_add_disk(self.h, _p(args[0]), _p(args[1]), height, radius, inner, material, None, 0)
        self.nprims += 1
        return self.nprims - 1

    def add_light(self, ltype, rgb, p=None, prim=-1, nsamples=1):
        L = np.asarray(rgb, np.float32)
        if ltype == 1:
            a, b, h, r, inner, mat = self.disks.pop(prim)
            self.lib.ref_pm_add_disk(self.h, _p(a), _p(b), h, r, inner, mat, _p(L), nsamples)
        else:
            pp = np.asarray(p, np.float32)
            self.lib.ref_pm_add_point_light(self.h, _p(pp), _p(L))

    def shoot(self, n_caustic, n_indirect, max_depth=5, ncores=1):
        """pbrt's PhotonIntegrator::Preprocess.  Returns (caustic, indirect) as (n,9) float arrays {p, alpha, wi},
        and (nCausticPaths, nIndirectPaths)."""
        if not self.built:
            self.lib.ref_pm_build(self.h)
            self.built = True
        cap_c, cap_i = 2 * n_caustic + 600000, 2 * n_indirect + 600000
        oc, oi = np.zeros((cap_c, 9), np.float32), np.zeros((cap_i, 9), np.float32)
        npaths, nst = (C.c_int64 * 2)(), (C.c_int64 * 2)()
        rc = self.lib.ref_pm_shoot(self.h, n_caustic, n_indirect, max_depth, ncores, _p(oc), cap_c, _p(oi), cap_i,
                                   npaths, nst)
        assert rc == 0
        return oc[: nst[0]].copy(), oi[: nst[1]].copy(), (npaths[0], npaths[1])

    def lphoton(self, photons9, n_paths, queries, k, max_dist2):
        ph = np.ascontiguousarray(photons9, np.float32).reshape(-1, 9)
        q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
        out = np.zeros((len(q), 3), np.float32)
        self.lib.ref_pm_lphoton(len(ph), _p(ph), int(n_paths), len(q), _p(q), k, max_dist2, _p(out))
        return out


def _ref_lphoton_timed(self, photons9, n_paths, queries, k, max_dist2, nthreads):
    """pbrt's KdTree build + LPhoton over all queries on nthreads threads: (rgb, kd-tree build seconds, lookup seconds)"""
    ph = np.ascontiguousarray(photons9, np.float32).reshape(-1, 9)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    out = np.zeros((len(q), 3), np.float32)
    b, l = C.c_double(), C.c_double()
    self.lib.ref_pm_lphoton_timed.argtypes = [C.c_int64, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_int, C.c_float, C.c_int,
                                              C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    self.lib.ref_pm_lphoton_timed.restype = None
    self.lib.ref_pm_lphoton_timed(len(ph), _p(ph), int(n_paths), len(q), _p(q), int(k), float(max_dist2), int(nthreads), _p(out),
                                  C.byref(b), C.byref(l))
    return out, b.value, l.value


def _ref_ephoton(self, photons9, n_paths, queries, k, max_dist2):
    """The reference's own EPhoton over its own kd-tree: (nq, 3)."""
    ph = np.ascontiguousarray(photons9, np.float32).reshape(-1, 9)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    out = np.zeros((len(q), 3), np.float32)
    self.lib.ref_pm_ephoton(len(ph), _p(ph), int(n_paths), len(q), _p(q), int(k), float(max_dist2), _p(out))
    return out


def _ref_radiance_lookup(self, photons9, queries):
    """The reference's own KdTree<RadiancePhoton>::Lookup + RadiancePhotonProcess; photons9 = {p, n, Lo}: Lo found (nq, 3)."""
    ph = np.ascontiguousarray(photons9, np.float32).reshape(-1, 9)
    q = np.ascontiguousarray(queries).view(np.float32).reshape(-1, 12)
    out = np.zeros((len(q), 3), np.float32)
    self.lib.ref_pm_radiance_lookup(len(ph), _p(ph), len(q), _p(q), _p(out))
    return out


RefPhotonScene.ephoton = _ref_ephoton
RefPhotonScene.lphoton_timed = _ref_lphoton_timed
RefPhotonScene.radiance_lookup = _ref_radiance_lookup


def oct_decode(w):
    """numpy restatement of the photon direction decoding (float32 throughout)."""
    w = np.asarray(w, np.uint32)
    x = (w & 0xFFFF).astype(np.uint16).view(np.int16).astype(np.float32) / np.float32(32767)
    y = (w >> 16).astype(np.uint16).view(np.int16).astype(np.float32) / np.float32(32767)
    z = np.float32(1) - np.abs(x) - np.abs(y)
    neg = z < 0
    ox = (np.float32(1) - np.abs(y)) * np.where(x >= 0, np.float32(1), np.float32(-1))
    oy = (np.float32(1) - np.abs(x)) * np.where(y >= 0, np.float32(1), np.float32(-1))
    x, y = np.where(neg, ox, x), np.where(neg, oy, y)
    v = np.stack([x, y, z], -1).astype(np.float32)
    return v / np.linalg.norm(v, axis=-1, keepdims=True).astype(np.float32)


def photons_to9(ph):
    out = np.zeros((len(ph), 9), np.float32)
    out[:, 0:3], out[:, 3:6], out[:, 6:9] = ph["p"], ph["alpha"], oct_decode(ph["wi_oct"])
    return out


# ---- film reconstruction (oracle_film.c; reference: pbrt's own ImageFilm through ref_film.cpp) -------------------------
FILTERS = {"box": 0, "triangle": 1, "gaussian": 2, "mitchell": 3, "sinc": 4}


def filter_args(name, xwidth=None, ywidth=None, **kw):
    """(type, xwidth, ywidth, p0, p1) with pbrt's defaults (filters/*.cpp Create*Filter)."""
    t = FILTERS[name]
    dw = {0: 0.5, 1: 2.0, 2: 2.0, 3: 2.0, 4: 4.0}[t]
    p0, p1 = {0: (0.0, 0.0), 1: (0.0, 0.0), 2: (kw.get("alpha", 2.0), 0.0), 3: (kw.get("B", 1 / 3), kw.get("C", 1 / 3)),
              4: (kw.get("tau", 3.0), 0.0)}[t]
    return t, float(dw if xwidth is None else xwidth), float(dw if ywidth is None else ywidth), float(p0), float(p1)


def filter_table(flt):
    lib = port()
    out = np.zeros(256, np.float32)
    lib.port_filter_table.argtypes = [C.c_int] + [C.c_float] * 4 + [C.c_void_p]
    lib.port_filter_table.restype = None
    lib.port_filter_table(flt[0], flt[1], flt[2], flt[3], flt[4], _p(out))
    return out


def film(flt, cells, spp, xy, rgb, pixels):
    """cells = (cx0, cy0, cx1, cy1) sample cells, pixels = (x0, y0, x1, y1); xy (n, 2), rgb (n, 3), n = cells * spp,
    cell row-major.  Returns (y1-y0, x1-x0, 3)."""
    lib = port()
    lib.port_film.argtypes = [C.c_int] + [C.c_float] * 4 + [C.c_int] * 5 + [C.c_void_p] * 2 + [C.c_int] * 4 + [C.c_void_p]
    lib.port_film.restype = None
    xy = np.ascontiguousarray(xy, np.float32)
    rgb = np.ascontiguousarray(rgb, np.float32)
    n = (cells[2] - cells[0]) * (cells[3] - cells[1]) * spp
    assert xy.shape == (n, 2) and rgb.shape == (n, 3)
    out = np.zeros((pixels[3] - pixels[1], pixels[2] - pixels[0], 3), np.float32)
    lib.port_film(flt[0], flt[1], flt[2], flt[3], flt[4], *[int(c) for c in cells], int(spp), _p(xy), _p(rgb),
                  *[int(c) for c in pixels], _p(out))
    return out


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = (int(v) for v in f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4").reshape(h, w, 3)
    return data[::-1].astype(np.float32)        # PFM rows run bottom to top


def ref_film(flt, xres, yres, xy, rgb, pixels=None):
    """pbrt's ImageFilm (oracle/_ref) fed with the samples in the given order; pixels = crop window in pixels."""
    import tempfile
    lib = C.CDLL(REF_SO)
    lib.ref_film.argtypes = [C.c_int] + [C.c_float] * 4 + [C.c_int] * 2 + [C.c_void_p, C.c_longlong, C.c_void_p, C.c_char_p]
    x0, y0, x1, y1 = pixels if pixels is not None else (0, 0, xres, yres)
    crop = np.array([x0 / xres, x1 / xres, y0 / yres, y1 / yres], np.float32)
    s = np.ascontiguousarray(np.concatenate([np.asarray(xy, np.float32), np.asarray(rgb, np.float32)], 1))
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "film.pfm")
        lib.ref_film(flt[0], flt[1], flt[2], flt[3], flt[4], xres, yres, _p(crop), len(s), _p(s), path.encode())
        return read_pfm(path)


# ---- image-plane sample patterns (oracle_film.c port_sample_positions; reference: pbrt's Sample02 / StratifiedSample2D)
SAMPLERS = {"random": 0, "stratified": 1, "lowdiscrepancy": 2}


def sample_positions(sampler, seed, xres, yres, window, spp, nx=1, ny=1):
    """window = (x0, y0, x1, y1) sample cells; returns (cells * spp, 2) raster positions, cell-major."""
    lib = port()
    lib.port_sample_positions.argtypes = [C.c_int] * 3 + [C.c_uint64] + [C.c_int] * 7 + [C.c_void_p]
    lib.port_sample_positions.restype = None
    n = (window[2] - window[0]) * (window[3] - window[1]) * spp
    out = np.zeros((n, 2), np.float32)
    lib.port_sample_positions(SAMPLERS[sampler], nx, ny, seed, xres, yres, *[int(v) for v in window], spp, _p(out))
    return out


def ld_point(n, s0, s1):
    lib = port()
    lib.port_vdc.restype = lib.port_sobol2.restype = C.c_float
    lib.port_vdc.argtypes = lib.port_sobol2.argtypes = [C.c_uint32, C.c_uint32]
    return np.float32(lib.port_vdc(n, s0)), np.float32(lib.port_sobol2(n, s1))


def ref_ld_point(n, s0, s1):
    lib = C.CDLL(REF_SO)
    lib.ref_sample02.argtypes = [C.c_uint, C.c_uint, C.c_uint, C.c_void_p]
    out = np.zeros(2, np.float32)
    lib.ref_sample02(n, s0, s1, _p(out))
    return out[0], out[1]


def ref_stratified_centres(nx, ny):
    lib = C.CDLL(REF_SO)
    lib.ref_stratified_centres.argtypes = [C.c_int, C.c_int, C.c_void_p]
    out = np.zeros((nx * ny, 2), np.float32)
    lib.ref_stratified_centres(nx, ny, _p(out))
    return out
