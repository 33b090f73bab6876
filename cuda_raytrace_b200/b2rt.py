"""ctypes binding of the b2rt C ABI (include/b2rt.h) -- used by tests/ and bench.py.

This is NOT the product's host layer (that is the C++ shim implementing the reference's
cuda_render/cudaapi.h:8-19 on top of the same C ABI); it only loads libb2rt.so and forwards.
There is no fallback of any kind: if the library is missing or no B200 is visible, calls raise.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B2RT_LIB") or os.path.join(_HERE, "libb2rt.so")     # B2RT_LIB: e.g. the -DB2RT_STATS build

HOST, DEVICE = 0, 1
SORT_NOALLOC_MAX = 1 << 20          # B2RT_SORT_NOALLOC_MAX (include/b2rt.h)
MAT_MATTE, MAT_MIRROR, MAT_GLASS = 0, 1, 2
LIGHT_POINT, LIGHT_DISK_AREA = 0, 1

RAY_DTYPE = np.dtype([("o", np.float32, 3), ("tmin", np.float32), ("d", np.float32, 3), ("tmax", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float32), ("prim", np.int32), ("b1", np.float32), ("b2", np.float32)])
SURFACE_DTYPE = np.dtype([("p", np.float32, 3), ("t", np.float32), ("ng", np.float32, 3), ("u", np.float32),
                          ("n", np.float32, 3), ("v", np.float32), ("sn", np.float32, 3), ("ray_eps", np.float32),
                          ("tn", np.float32, 3), ("prim", np.int32)])      # == b2rt_surface
MESH_REVERSE = 1


class AccelInfo(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("n_tris", C.c_int32), ("n_wide_nodes", C.c_int32),
                ("n_leaf_slots", C.c_int32), ("max_depth", C.c_int32), ("wide_width", C.c_int32),
                ("bounds", C.c_float * 6), ("sah_cost", C.c_float), ("build_ms", C.c_float), ("upload_ms", C.c_float)]


class B2rtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b2rt error {code}: {msg}")
        self.code = code


# every symbol include/b2rt.h declares (tests check the .so exports exactly these)
SYMBOLS = [
    "b2rt_create", "b2rt_destroy", "b2rt_last_error", "b2rt_device_count", "b2rt_add_mesh", "b2rt_add_sphere",
    "b2rt_add_disk", "b2rt_add_material", "b2rt_add_light", "b2rt_clear_scene", "b2rt_build_accel",
    "b2rt_accel_info_get", "b2rt_accel_download", "b2rt_trace_closest", "b2rt_trace_any", "b2rt_sort_pairs", "b2rt_sort_pairs_noalloc",
    "b2rt_add_mesh_ex", "b2rt_add_sphere_ex", "b2rt_add_disk_ex", "b2rt_surface_at", "b2rt_scene_save", "b2rt_scene_load", "b2rt_scene_counts", "b2rt_timings_enable", "b2rt_get_timings", "b2rt_object_begin", "b2rt_object_end", "b2rt_add_instance", "b2rt_instance_info", "b2rt_trace_stats", "b2rt_add_spot_light", "b2rt_add_distant_light", "b2rt_sync", "b2rt_timer_start", "b2rt_timer_stop", "b2rt_launch_count", "b2rt_stream", "b2rt_host_alloc", "b2rt_host_free",
]

_lib = None


def load():
    """Load libb2rt.so (built by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, fp = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_float)
    ip = C.POINTER(C.c_int32)
    lib.b2rt_create.argtypes = [C.c_int, C.c_uint64, C.POINTER(vp)]
    lib.b2rt_destroy.argtypes = [vp]
    lib.b2rt_destroy.restype = None
    lib.b2rt_last_error.argtypes = [vp]
    lib.b2rt_last_error.restype = C.c_char_p
    lib.b2rt_device_count.argtypes = []
    lib.b2rt_add_mesh.argtypes = [vp, i32, vp, i32, vp, vp, vp, i32, i32, ip]
    lib.b2rt_add_sphere.argtypes = [vp, vp, vp, C.c_float, i32, i32, ip]
    lib.b2rt_add_disk.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_float, i32, i32, ip]
    lib.b2rt_add_material.argtypes = [vp, i32, vp, vp, C.c_float, ip]
    lib.b2rt_add_light.argtypes = [vp, i32, vp, vp, i32, i32, ip]
    lib.b2rt_clear_scene.argtypes = [vp]
    lib.b2rt_build_accel.argtypes = [vp]
    lib.b2rt_accel_info_get.argtypes = [vp, C.POINTER(AccelInfo)]
    lib.b2rt_accel_download.argtypes = [vp] + [vp] * 7
    lib.b2rt_trace_closest.argtypes = [vp, i64, vp, vp, C.c_int]
    lib.b2rt_add_mesh_ex.argtypes = [vp, i32, vp, i32, vp, vp, vp, i32, i32, i32, ip]
    lib.b2rt_surface_at.argtypes = [vp, i64, vp, vp, C.c_int]
    lib.b2rt_trace_any.argtypes = [vp, i64, vp, vp, C.c_int]
    lib.b2rt_sort_pairs.argtypes = [vp, i64, vp, vp, C.c_int]
    lib.b2rt_sort_pairs_noalloc.argtypes = [vp, i64, vp, vp, C.c_int]
    lib.b2rt_sync.argtypes = [vp]
    lib.b2rt_timer_start.argtypes = [vp]
    lib.b2rt_timer_stop.argtypes = [vp, fp]
    lib.b2rt_launch_count.argtypes = [vp, C.POINTER(i64), C.c_int]
    lib.b2rt_stream.argtypes = [vp]
    lib.b2rt_stream.restype = vp
    lib.b2rt_host_alloc.argtypes = [vp, i64]
    lib.b2rt_host_alloc.restype = vp
    lib.b2rt_host_free.argtypes = [vp, vp]
    lib.b2rt_host_free.restype = None
    _lib = lib
    return lib


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


class Context:
    """One b2rt context = one GPU (the reference's CudaRender singleton, cuda_render/cudaapi.cpp:3)."""

    def __init__(self, device=0, seed=0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.b2rt_create(device, seed, C.byref(h))
        if rc:
            raise B2rtError(rc, self.lib.b2rt_last_error(None).decode())
        self.h = h

    def close(self):
        if getattr(self, "h", None):
            for ptr in getattr(self, "_pinned", []):
                self.lib.b2rt_host_free(self.h, ptr)
            self._pinned = []
            self.lib.b2rt_destroy(self.h)
            self.h = None

    def pinned_empty(self, shape, dtype=np.float32):
        """numpy array in page-locked host memory (b2rt_host_alloc), owned by the context and valid until close():
        the buffer to hand to the host-pointer calls (`out=` of render_frame) so the copies run at PCIe speed."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        ptr = self.lib.b2rt_host_alloc(self.h, max(n, 1))
        if not ptr:
            raise B2rtError(-2, self.lib.b2rt_last_error(self.h).decode())
        if not hasattr(self, "_pinned"):
            self._pinned = []
        self._pinned.append(ptr)
        buf = (C.c_char * max(n, 1)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc:
            raise B2rtError(rc, self.lib.b2rt_last_error(self.h).decode())

    def trace_stats(self, reset=True):
        """(wide nodes fetched, primitive records tested) by the traversal kernel; needs the -DB2RT_STATS library
        (B2RT_LIB=.../libb2rt_stats.so)"""
        a, b = C.c_uint64(), C.c_uint64()
        self.lib.b2rt_trace_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_int]
        self._ck(self.lib.b2rt_trace_stats(self.h, C.byref(a), C.byref(b), 1 if reset else 0))
        return a.value, b.value

    def last_error(self):
        """message of the last failing or warning call (e.g. b2rt_photons_shoot giving up like pbrt does)"""
        return self.lib.b2rt_last_error(self.h).decode()

    # ---- scene --------------------------------------------------------------------------------
    def add_mesh(self, P, idx, material=-1, arealight=-1, N=None, uv=None, reverse=False):
        """N: per-vertex normals in WORLD space, uv: per-vertex (u, v) -- pbrt's "N" / "uv" mesh parameters;
        reverse: pbrt ReverseOrientation ^ TransformSwapsHandedness."""
        P = _f32(P, (-1, 3))
        idx = np.ascontiguousarray(idx, dtype=np.int32).reshape(-1, 3)
        N = None if N is None else _f32(N, (-1, 3))
        uv = None if uv is None else _f32(uv, (-1, 2))
        if (N is not None and len(N) != len(P)) or (uv is not None and len(uv) != len(P)):
            raise ValueError("add_mesh: N / uv need one entry per vertex")
        first = C.c_int32()
        self._ck(self.lib.b2rt_add_mesh_ex(self.h, len(P), _ptr(P), len(idx), _ptr(idx), _ptr(N), _ptr(uv), material,
                                           arealight, MESH_REVERSE if reverse else 0, C.byref(first)))
        return first.value

    def surface_at(self, rays):
        """closest hit + the hit attributes the shade stage derives (b2rt_surface_at): SURFACE_DTYPE per ray"""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        out = np.zeros(len(rays), SURFACE_DTYPE)
        self._ck(self.lib.b2rt_surface_at(self.h, len(rays), _ptr(rays), _ptr(out), HOST))
        return out

    def add_sphere(self, o2w, radius, w2o=None, material=-1, arealight=-1, zmin=None, zmax=None, phimax=None):
        """zmin / zmax / phimax (degrees): pbrt's partial sphere"""
        o2w = _f32(o2w, (12,))
        w2o = None if w2o is None else _f32(w2o, (12,))
        p = C.c_int32()
        if zmin is not None or zmax is not None or phimax is not None:
            self.lib.b2rt_add_sphere_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float,
                                                    C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
            self._ck(self.lib.b2rt_add_sphere_ex(self.h, _ptr(o2w), _ptr(w2o), radius, -radius if zmin is None else zmin,
                                                 radius if zmax is None else zmax, 360.0 if phimax is None else phimax,
                                                 material, arealight, C.byref(p)))
            return p.value
        self._ck(self.lib.b2rt_add_sphere(self.h, _ptr(o2w), _ptr(w2o), radius, material, arealight, C.byref(p)))
        return p.value

    def add_disk(self, o2w, height, radius, inner=0.0, w2o=None, material=-1, arealight=-1, phimax=None):
        o2w = _f32(o2w, (12,))
        w2o = None if w2o is None else _f32(w2o, (12,))
        p = C.c_int32()
        if phimax is not None:
            self.lib.b2rt_add_disk_ex.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_float,
                                                  C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
            self._ck(self.lib.b2rt_add_disk_ex(self.h, _ptr(o2w), _ptr(w2o), height, radius, inner, phimax, material, arealight, C.byref(p)))
            return p.value
        self._ck(self.lib.b2rt_add_disk(self.h, _ptr(o2w), _ptr(w2o), height, radius, inner, material, arealight,
                                        C.byref(p)))
        return p.value

    def add_material(self, mtype, rgb, kt=None, ior=1.5):
        rgb = _f32(rgb, (3,))
        kt = None if kt is None else _f32(kt, (3,))
        m = C.c_int32()
        self._ck(self.lib.b2rt_add_material(self.h, mtype, _ptr(rgb), _ptr(kt), ior, C.byref(m)))
        return m.value

    def add_light(self, ltype, rgb, p=None, prim=-1, nsamples=1):
        rgb = _f32(rgb, (3,))
        p = None if p is None else _f32(p, (3,))
        l = C.c_int32()
        self._ck(self.lib.b2rt_add_light(self.h, ltype, _ptr(p), _ptr(rgb), prim, nsamples, C.byref(l)))
        return l.value

    def add_spot_light(self, l2w, w2l, rgb, cos_total, cos_falloff):
        """pbrt SpotLight: LightToWorld / WorldToLight (3x4 row-major, +z = axis), I, cos(coneangle), cos(coneangle - delta)"""
        l2w, w2l, rgb = _f32(l2w, (12,)), _f32(w2l, (12,)), _f32(rgb, (3,))
        l = C.c_int32()
        self.lib.b2rt_add_spot_light.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.POINTER(C.c_int32)]
        self._ck(self.lib.b2rt_add_spot_light(self.h, _ptr(l2w), _ptr(w2l), _ptr(rgb), cos_total, cos_falloff, C.byref(l)))
        return l.value

    def add_distant_light(self, dir_to_light, rgb):
        """pbrt DistantLight: unit direction TO the light, radiance L"""
        d, rgb = _f32(dir_to_light, (3,)), _f32(rgb, (3,))
        l = C.c_int32()
        self.lib.b2rt_add_distant_light.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]
        self._ck(self.lib.b2rt_add_distant_light(self.h, _ptr(d), _ptr(rgb), C.byref(l)))
        return l.value

    def clear_scene(self):
        self._ck(self.lib.b2rt_clear_scene(self.h))

    # ---- per-stage device times ------------------------------------------------------------------------
    STAGES = ("build", "trace_closest", "trace_any", "photons", "grid", "gather", "render", "film")

    def timings_enable(self, on=True):
        self.lib.b2rt_timings_enable.argtypes = [C.c_void_p, C.c_int]
        self._ck(self.lib.b2rt_timings_enable(self.h, 1 if on else 0))

    def timings(self, reset=False):
        """{stage: (device ms, calls)} accumulated while timings were enabled (b2rt_get_timings)"""
        class T(C.Structure):
            _fields_ = [("ms", C.c_double * 8), ("calls", C.c_int64 * 8)]
        t = T()
        self.lib.b2rt_get_timings.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        self._ck(self.lib.b2rt_get_timings(self.h, C.byref(t), 1 if reset else 0))
        return {n: (t.ms[i], t.calls[i]) for i, n in enumerate(self.STAGES)}

    # ---- binary scene cache ---------------------------------------------------------------------------
    def scene_save(self, path):
        self.lib.b2rt_scene_save.argtypes = [C.c_void_p, C.c_char_p]
        self._ck(self.lib.b2rt_scene_save(self.h, str(path).encode()))

    def scene_load(self, path):
        self.lib.b2rt_scene_load.argtypes = [C.c_void_p, C.c_char_p]
        self._ck(self.lib.b2rt_scene_load(self.h, str(path).encode()))

    def scene_counts(self):
        """{items, lights, materials, objects, instances} staged in the context"""
        c = (C.c_int32 * 5)()
        self.lib.b2rt_scene_counts.argtypes = [C.c_void_p, C.c_void_p]
        self._ck(self.lib.b2rt_scene_counts(self.h, c))
        return dict(zip(("items", "lights", "materials", "objects", "instances"), list(c)))

    # ---- objects / instances (two-level BVH) ------------------------------------------------------
    def object_begin(self):
        """the add_mesh / add_sphere / add_disk calls until object_end() define an object (ids local to it)"""
        o = C.c_int32()
        self.lib.b2rt_object_begin.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        self._ck(self.lib.b2rt_object_begin(self.h, C.byref(o)))
        return o.value

    def object_end(self):
        self.lib.b2rt_object_end.argtypes = [C.c_void_p]
        self._ck(self.lib.b2rt_object_end(self.h))

    def add_instance(self, obj, o2w, w2o=None):
        """place object `obj` under the instance-to-world transform o2w (3x4 row-major); returns the instance id"""
        o2w = _f32(o2w, (12,))
        w2o = None if w2o is None else _f32(w2o, (12,))
        k = C.c_int32()
        self.lib.b2rt_add_instance.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.POINTER(C.c_int32)]
        self._ck(self.lib.b2rt_add_instance(self.h, obj, _ptr(o2w), _ptr(w2o), C.byref(k)))
        return k.value

    def instance_info(self, k):
        """(first global primitive id, primitive count, object id) of instance k; after build_accel"""
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        self.lib.b2rt_instance_info.argtypes = [C.c_void_p, C.c_int32] + [C.POINTER(C.c_int32)] * 3
        self._ck(self.lib.b2rt_instance_info(self.h, k, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    # ---- accel --------------------------------------------------------------------------------
    def build_accel(self):
        self._ck(self.lib.b2rt_build_accel(self.h))
        return self.accel_info()

    def accel_info(self):
        info = AccelInfo()
        self._ck(self.lib.b2rt_accel_info_get(self.h, C.byref(info)))
        return info

    def accel_download(self):
        info = self.accel_info()
        n, nw = info.n_prims, info.n_wide_nodes
        out = {
            "morton": np.zeros(n, np.uint32), "prim_order": np.zeros(n, np.int32),
            "bin_parent": np.zeros(max(2 * n - 1, 0), np.int32), "bin_child": np.zeros(max(2 * (n - 1), 0), np.int32),
            "bin_box": np.zeros((max(2 * n - 1, 0), 6), np.float32), "wide_nodes": np.zeros((nw, 32), np.uint32),
            "leaf_prims": np.zeros(n, np.int32),
        }
        self._ck(self.lib.b2rt_accel_download(self.h, *[_ptr(out[k]) for k in (
            "morton", "prim_order", "bin_parent", "bin_child", "bin_box", "wide_nodes", "leaf_prims")]))
        return out

    # ---- rays ---------------------------------------------------------------------------------
    def trace_closest(self, rays):
        """Host-buffer call: rays is an (n,8) float32 / RAY_DTYPE array; returns a HIT_DTYPE array."""
        rays = np.ascontiguousarray(rays)
        n = rays.shape[0]
        hits = np.zeros(n, HIT_DTYPE)
        self._ck(self.lib.b2rt_trace_closest(self.h, n, _ptr(rays), _ptr(hits), HOST))
        return hits

    def trace_any(self, rays):
        rays = np.ascontiguousarray(rays)
        n = rays.shape[0]
        occ = np.zeros(n, np.uint8)
        self._ck(self.lib.b2rt_trace_any(self.h, n, _ptr(rays), _ptr(occ), HOST))
        return occ

    def trace_closest_device(self, nrays, rays_ptr, hits_ptr):
        self._ck(self.lib.b2rt_trace_closest(self.h, nrays, C.c_void_p(rays_ptr), C.c_void_p(hits_ptr), DEVICE))

    def trace_any_device(self, nrays, rays_ptr, occ_ptr):
        self._ck(self.lib.b2rt_trace_any(self.h, nrays, C.c_void_p(rays_ptr), C.c_void_p(occ_ptr), DEVICE))

    def sort_pairs_device(self, n, keys_ptr, vals_ptr, key_bits=32):
        self._ck(self.lib.b2rt_sort_pairs(self.h, n, C.c_void_p(keys_ptr), C.c_void_p(vals_ptr), key_bits))

    def sort_pairs_noalloc_device(self, n, keys_ptr, vals_ptr, key_bits=32):
        """b2rt_sort_pairs_noalloc: scratch in the library's static device array, n <= SORT_NOALLOC_MAX, no cudaMalloc."""
        self._ck(self.lib.b2rt_sort_pairs_noalloc(self.h, n, C.c_void_p(keys_ptr), C.c_void_p(vals_ptr), key_bits))

    # ---- timing -------------------------------------------------------------------------------
    def sync(self):
        self._ck(self.lib.b2rt_sync(self.h))

    def timer_start(self):
        self._ck(self.lib.b2rt_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float()
        self._ck(self.lib.b2rt_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def launch_count(self, reset=False):
        n = C.c_int64()
        self._ck(self.lib.b2rt_launch_count(self.h, C.byref(n), int(reset)))
        return n.value

    @property
    def stream(self):
        return self.lib.b2rt_stream(self.h)


# ---- photon pass / hash grid / radiance estimate (include/b2rt.h, second half) ---------------------------------------
MAP_CAUSTIC, MAP_INDIRECT = 0, 1
PHOTON_DTYPE = np.dtype([("p", np.float32, 3), ("wi_oct", np.uint32), ("alpha", np.float32, 3), ("path", np.uint32)])
QUERY_DTYPE = np.dtype([("p", np.float32, 3), ("n", np.float32, 3), ("kd", np.float32, 3), ("pad", np.float32, 3)])


class GridInfo(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("cell_size", C.c_float), ("inv_cell", C.c_float), ("dims", C.c_int32 * 3),
                ("n_cells", C.c_int32), ("n_photons", C.c_int64), ("build_ms", C.c_float)]


SYMBOLS += [
    "b2rt_photons_trace", "b2rt_photons_shoot", "b2rt_photons_clear", "b2rt_photons_count", "b2rt_photons_download",
    "b2rt_photons_upload", "b2rt_photons_device_ptr", "b2rt_photon_grid_build", "b2rt_photon_grid_info",
    "b2rt_photon_grid_download", "b2rt_gather", "b2rt_gather_sets",
]


def _bind_photon(lib):
    vp, i32, i64, f = C.c_void_p, C.c_int32, C.c_int64, C.c_float
    lib.b2rt_photons_trace.argtypes = [vp, i64, i64, i32, i32, vp]
    lib.b2rt_photons_shoot.argtypes = [vp, i64, i64, i32, i64, vp, vp]
    lib.b2rt_photons_clear.argtypes = [vp]
    lib.b2rt_photons_count.argtypes = [vp, i32, C.POINTER(i64), C.POINTER(i64)]
    lib.b2rt_photons_download.argtypes = [vp, i32, vp, i64, C.c_int]
    lib.b2rt_photons_upload.argtypes = [vp, i32, vp, i64, i64, C.c_int]
    lib.b2rt_photons_device_ptr.argtypes = [vp, i32]
    lib.b2rt_photons_device_ptr.restype = vp
    lib.b2rt_photon_grid_build.argtypes = [vp, i32, f, f]
    lib.b2rt_photon_grid_info.argtypes = [vp, i32, C.POINTER(GridInfo)]
    lib.b2rt_photon_grid_download.argtypes = [vp, i32, vp, vp, vp, vp]
    lib.b2rt_gather.argtypes = [vp, i32, i64, vp, i32, f, vp, vp, vp, C.c_int]
    lib.b2rt_gather_sets.argtypes = [vp, i32, i64, vp, i32, f, vp, C.c_int]


_load_core = load


def load():  # noqa: F811
    lib = _load_core()
    if not getattr(lib, "_b2rt_photon_bound", False):
        _bind_photon(lib)
        lib._b2rt_photon_bound = True
    return lib


def _photons_trace(self, first_path, n_paths, max_depth=5, caustic=True, indirect=True):
    st = (C.c_int64 * 2)()
    self._ck(self.lib.b2rt_photons_trace(self.h, first_path, n_paths, max_depth, (1 if caustic else 0) | (2 if indirect else 0), st))
    return st[0], st[1]


def _photons_shoot(self, n_caustic, n_indirect, max_depth=5, batch_paths=65536):
    npaths, nst = (C.c_int64 * 2)(), (C.c_int64 * 2)()
    self._ck(self.lib.b2rt_photons_shoot(self.h, n_caustic, n_indirect, max_depth, batch_paths, npaths, nst))
    return (npaths[0], npaths[1]), (nst[0], nst[1])


def _photons_clear(self):
    self._ck(self.lib.b2rt_photons_clear(self.h))


def _photons_count(self, which):
    n, p = C.c_int64(), C.c_int64()
    self._ck(self.lib.b2rt_photons_count(self.h, which, C.byref(n), C.byref(p)))
    return n.value, p.value


def _photons_download(self, which):
    n, _ = self.photons_count(which)
    out = np.zeros(n, PHOTON_DTYPE)
    self._ck(self.lib.b2rt_photons_download(self.h, which, _ptr(out), n, HOST))
    return out


def _photons_upload(self, which, photons, n_paths):
    photons = np.ascontiguousarray(photons, PHOTON_DTYPE)
    self._ck(self.lib.b2rt_photons_upload(self.h, which, _ptr(photons), len(photons), n_paths, HOST))


def _grid_build(self, which, cell_size=0.0, target_per_cell=0.0):
    self._ck(self.lib.b2rt_photon_grid_build(self.h, which, cell_size, target_per_cell))
    gi = GridInfo()
    self._ck(self.lib.b2rt_photon_grid_info(self.h, which, C.byref(gi)))
    return gi


def _grid_info(self, which):
    gi = GridInfo()
    self._ck(self.lib.b2rt_photon_grid_info(self.h, which, C.byref(gi)))
    return gi


def _grid_download(self, which):
    gi = GridInfo()
    self._ck(self.lib.b2rt_photon_grid_info(self.h, which, C.byref(gi)))
    n = gi.n_photons
    out = {"keys": np.zeros(n, np.uint32), "order": np.zeros(n, np.uint32),
           "cell_start": np.zeros(gi.n_cells + 1, np.uint32), "sorted": np.zeros(n, PHOTON_DTYPE), "info": gi}
    self._ck(self.lib.b2rt_photon_grid_download(self.h, which, _ptr(out["keys"]), _ptr(out["order"]),
                                                _ptr(out["cell_start"]), _ptr(out["sorted"])))
    return out


def make_queries(p, n, kd):
    q = np.zeros(len(p), QUERY_DTYPE)
    q["p"], q["n"], q["kd"] = p, n, kd
    return q


def _gather(self, which, queries, k, max_dist2):
    queries = np.ascontiguousarray(queries, QUERY_DTYPE)
    nq = len(queries)
    rgb, r2, found = np.zeros((nq, 3), np.float32), np.zeros(nq, np.float32), np.zeros(nq, np.int32)
    self._ck(self.lib.b2rt_gather(self.h, which, nq, _ptr(queries), k, max_dist2, _ptr(rgb), _ptr(r2), _ptr(found), HOST))
    return rgb, r2, found


def _gather_sets(self, which, queries, k, max_dist2):
    queries = np.ascontiguousarray(queries, QUERY_DTYPE)
    idx = np.zeros((len(queries), k), np.int32)
    self._ck(self.lib.b2rt_gather_sets(self.h, which, len(queries), _ptr(queries), k, max_dist2, _ptr(idx), HOST))
    return idx


for _n, _f in [("photons_trace", _photons_trace), ("photons_shoot", _photons_shoot), ("photons_clear", _photons_clear),
               ("photons_count", _photons_count), ("photons_download", _photons_download),
               ("photons_upload", _photons_upload), ("grid_build", _grid_build), ("grid_download", _grid_download),
               ("gather", _gather), ("gather_sets", _gather_sets), ("grid_info", _grid_info)]:
    setattr(Context, _n, _f)


# ---- camera pass ------------------------------------------------------------------------------------------------------
class Camera(C.Structure):
    _fields_ = [("o", C.c_float * 3), ("A", C.c_float * 3), ("B", C.c_float * 3), ("C", C.c_float * 3),
                ("xres", C.c_int32), ("yres", C.c_int32)]


class FrameParams(C.Structure):
    _fields_ = [("spp", C.c_int32), ("max_specular_depth", C.c_int32), ("nused", C.c_int32), ("max_dist2", C.c_float),
                ("use_caustic", C.c_int32), ("use_indirect", C.c_int32), ("direct_lighting", C.c_int32),
                ("x0", C.c_int32), ("x1", C.c_int32), ("y0", C.c_int32), ("y1", C.c_int32)]


SYMBOLS += ["b2rt_render_frame", "b2rt_camera_samples"]


def pbrt_camera(eye, look, up, fov, xres, yres):
    """pbrt `LookAt` + `Camera "perspective"` (core/transform.cpp LookAt, cameras/perspective.cpp:45-78): left-handed,
    fov spans the shorter image axis.  Returns the b2rt_camera whose rays equal GenerateRay's for lensradius 0."""
    eye, look, up = (np.asarray(v, np.float64) for v in (eye, look, up))
    d = look - eye
    d /= np.linalg.norm(d)
    left = np.cross(up / np.linalg.norm(up), d)
    left /= np.linalg.norm(left)
    nup = np.cross(d, left)
    aspect = xres / yres
    sw = (-aspect, aspect, -1.0, 1.0) if aspect > 1 else (-1.0, 1.0, -1.0 / aspect, 1.0 / aspect)
    t = np.tan(np.radians(fov) / 2)
    cam = Camera()
    A = sw[0] * t * left + sw[3] * t * nup + d
    B = (sw[1] - sw[0]) / xres * t * left
    Cv = -(sw[3] - sw[2]) / yres * t * nup
    for dst, src in ((cam.o, eye), (cam.A, A), (cam.B, B), (cam.C, Cv)):
        for i in range(3):
            dst[i] = float(src[i])
    cam.xres, cam.yres = xres, yres
    return cam


def _render_frame(self, cam, spp=1, max_specular_depth=5, nused=50, max_dist=0.1, caustic=True, indirect=True,
                  direct=True, window=None, out=None):
    """out: optional (y1-y0, x1-x0, 3) float32 array to receive the image (e.g. from pinned_empty); default a new array."""
    x0, x1, y0, y1 = window if window else (0, cam.xres, 0, cam.yres)
    fp = FrameParams(spp, max_specular_depth, nused, float(np.float32(max_dist) * np.float32(max_dist)), int(caustic),
                     int(indirect), int(direct), x0, x1, y0, y1)
    if out is None:
        out = np.zeros((y1 - y0, x1 - x0, 3), np.float32)
    elif out.shape != (y1 - y0, x1 - x0, 3) or out.dtype != np.float32 or not out.flags["C_CONTIGUOUS"]:
        raise ValueError("render_frame: out must be a C-contiguous float32 array of shape (y1-y0, x1-x0, 3)")
    self.lib.b2rt_render_frame.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(FrameParams), C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_render_frame(self.h, C.byref(cam), C.byref(fp), _ptr(out), HOST))
    return out


def _render_frame_device(self, cam, out_ptr, spp=1, max_specular_depth=5, nused=50, max_dist=0.1, caustic=True,
                         indirect=True, direct=True, window=None):
    """Same as render_frame, but the (y1-y0, x1-x0, 3) float32 image is written to the DEVICE pointer `out_ptr`
    (b2rt_render_frame(..., B2RT_DEVICE)); the call returns after the library stream has finished."""
    x0, x1, y0, y1 = window if window else (0, cam.xres, 0, cam.yres)
    fp = FrameParams(spp, max_specular_depth, nused, float(np.float32(max_dist) * np.float32(max_dist)), int(caustic),
                     int(indirect), int(direct), x0, x1, y0, y1)
    self.lib.b2rt_render_frame.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(FrameParams), C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_render_frame(self.h, C.byref(cam), C.byref(fp), C.c_void_p(out_ptr), DEVICE))


def _camera_samples(self, cam, spp=1, max_specular_depth=5, direct=True, window=None, queries=False):
    """b2rt_camera_samples: (rgb (n, 3), leaf counts (n,), leaf queries (n, slots) or None) of the crop window's samples"""
    x0, x1, y0, y1 = window if window else (0, cam.xres, 0, cam.yres)
    fp = FrameParams(spp, max_specular_depth, 50, 0.01, 0, 0, int(direct), x0, x1, y0, y1)
    n = (x1 - x0) * (y1 - y0) * spp
    rgb, nq, slots = np.zeros((n, 3), np.float32), np.zeros(n, np.uint32), C.c_int32()
    self.lib.b2rt_camera_samples.argtypes = [C.c_void_p, C.POINTER(Camera), C.POINTER(FrameParams), C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.POINTER(C.c_int32)]
    q = None
    if queries:
        self._ck(self.lib.b2rt_camera_samples(self.h, C.byref(cam), C.byref(fp), _ptr(rgb), _ptr(nq), None, C.byref(slots)))
        q = np.zeros((n, slots.value), QUERY_DTYPE)
    self._ck(self.lib.b2rt_camera_samples(self.h, C.byref(cam), C.byref(fp), _ptr(rgb), _ptr(nq), _ptr(q), C.byref(slots)))
    return rgb, nq, q


Context.camera_samples = _camera_samples
Context.render_frame = _render_frame
Context.render_frame_device = _render_frame_device


# ---- film reconstruction with pbrt's pixel filters (csrc/film.cu) ------------------------------------------------------
SYMBOLS += ["b2rt_film_set_filter", "b2rt_film_reconstruct"]
FILTERS = {"box": 0, "triangle": 1, "gaussian": 2, "mitchell": 3, "sinc": 4}


class Filter(C.Structure):
    _fields_ = [("type", C.c_int32), ("xwidth", C.c_float), ("ywidth", C.c_float), ("p0", C.c_float), ("p1", C.c_float)]


def pbrt_filter(name, xwidth=None, ywidth=None, **params):
    """pbrt `PixelFilter "name"` with its parameter names and defaults (filters/*.cpp Create*Filter): box 0.5,
    triangle 2, gaussian 2 (alpha 2), mitchell 2 (B = C = 1/3), sinc 4 (tau 3)."""
    t = FILTERS[name]
    dw = (0.5, 2.0, 2.0, 2.0, 4.0)[t]
    known = {0: (), 1: (), 2: ("alpha",), 3: ("B", "C"), 4: ("tau",)}[t]
    bad = set(params) - set(known)
    if bad:
        raise ValueError("filter %r has no parameter %s" % (name, sorted(bad)))
    p0 = {2: params.get("alpha", 2.0), 3: params.get("B", 1.0 / 3.0), 4: params.get("tau", 3.0)}.get(t, 0.0)
    p1 = params.get("C", 1.0 / 3.0) if t == 3 else 0.0
    return Filter(t, dw if xwidth is None else xwidth, dw if ywidth is None else ywidth, p0, p1)


def _set_filter(self, flt=None):
    """Pixel filter of the following render_frame calls (a Filter from pbrt_filter); None = pbrt's default box 0.5."""
    self.lib.b2rt_film_set_filter.argtypes = [C.c_void_p, C.POINTER(Filter)]
    self._ck(self.lib.b2rt_film_set_filter(self.h, C.byref(flt) if flt is not None else None))


def _film_reconstruct(self, flt, cells, spp, xy, rgb, pixels):
    """cells = (cx0, cy0, cx1, cy1), pixels = (x0, y0, x1, y1); xy (n, 2) / rgb (n, 3) float32, n = cells * spp in cell
    row-major order.  Returns the (y1-y0, x1-x0, 3) image."""
    xy = np.ascontiguousarray(xy, np.float32)
    rgb = np.ascontiguousarray(rgb, np.float32)
    n = (cells[2] - cells[0]) * (cells[3] - cells[1]) * spp
    if xy.shape != (n, 2) or rgb.shape != (n, 3):
        raise ValueError("film_reconstruct: xy / rgb must hold cells * spp samples")
    out = np.zeros((pixels[3] - pixels[1], pixels[2] - pixels[0], 3), np.float32)
    c4, p4 = (C.c_int32 * 4)(*[int(v) for v in cells]), (C.c_int32 * 4)(*[int(v) for v in pixels])
    self.lib.b2rt_film_reconstruct.argtypes = [C.c_void_p, C.POINTER(Filter), C.POINTER(C.c_int32), C.c_int32, C.c_void_p,
                                               C.c_void_p, C.POINTER(C.c_int32), C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_film_reconstruct(self.h, C.byref(flt), c4, int(spp), _ptr(xy), _ptr(rgb), p4, _ptr(out), HOST))
    return out


def _film_reconstruct_device(self, flt, cells, spp, xy_ptr, rgb_ptr, pixels, out_ptr):
    """film_reconstruct on DEVICE pointers (asynchronous on the library stream)."""
    c4, p4 = (C.c_int32 * 4)(*[int(v) for v in cells]), (C.c_int32 * 4)(*[int(v) for v in pixels])
    self.lib.b2rt_film_reconstruct.argtypes = [C.c_void_p, C.POINTER(Filter), C.POINTER(C.c_int32), C.c_int32, C.c_void_p,
                                               C.c_void_p, C.POINTER(C.c_int32), C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_film_reconstruct(self.h, C.byref(flt), c4, int(spp), C.c_void_p(xy_ptr), C.c_void_p(rgb_ptr), p4,
                                            C.c_void_p(out_ptr), DEVICE))


Context.set_filter = _set_filter
Context.film_reconstruct = _film_reconstruct
Context.film_reconstruct_device = _film_reconstruct_device


# ---- image-plane sample patterns ----------------------------------------------------------------------------------------
SYMBOLS += ["b2rt_sampler_set", "b2rt_sample_positions"]
SAMPLERS = {"random": 0, "stratified": 1, "lowdiscrepancy": 2}


def _set_sampler(self, name="random", xsamples=1, ysamples=1):
    """pbrt `Sampler "name"`: random (default), stratified (xsamples * ysamples must equal spp) or lowdiscrepancy."""
    self.lib.b2rt_sampler_set.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_int32]
    self._ck(self.lib.b2rt_sampler_set(self.h, SAMPLERS[name] if isinstance(name, str) else int(name), int(xsamples), int(ysamples)))


def _sample_positions(self, xres, yres, window, spp):
    """(cells * spp, 2) raster positions of the camera samples of window = (x0, y0, x1, y1), cell-major."""
    n = (window[2] - window[0]) * (window[3] - window[1]) * spp
    out = np.zeros((max(n, 0), 2), np.float32)
    w4 = (C.c_int32 * 4)(*[int(v) for v in window])
    self.lib.b2rt_sample_positions.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.c_void_p,
                                               C.c_int]
    self._ck(self.lib.b2rt_sample_positions(self.h, int(xres), int(yres), w4, int(spp), _ptr(out), HOST))
    return out


Context.set_sampler = _set_sampler
Context.sample_positions = _sample_positions


# ---- final gathering (pbrt photonmap with finalgather true) ------------------------------------------
SYMBOLS += ["b2rt_fg_deposit_limits", "b2rt_multi_fg_deposit_limits", "b2rt_photons_trace_fg", "b2rt_irradiance", "b2rt_radiance_compute", "b2rt_radiance_lookup",
            "b2rt_final_gather_set", "b2rt_final_gather"]
MAP_CAUSTIC, MAP_INDIRECT, MAP_DIRECT, MAP_RADIANCE = 0, 1, 2, 3


def _photons_trace_fg(self, first_path, n_paths, max_depth=5):
    """Direct photons (map 2) and radiance photons (map 3) of paths [first_path, first_path + n_paths)."""
    st = (C.c_int64 * 2)()
    self.lib.b2rt_photons_trace_fg.argtypes = [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_void_p]
    self._ck(self.lib.b2rt_photons_trace_fg(self.h, int(first_path), int(n_paths), int(max_depth), st))
    return st[0], st[1]


def _fg_deposit_limits(self, caustic_paths=-1, indirect_paths=-1):
    """pbrt's depositedPhoton rule for the final-gathering photon pass (b2rt_fg_deposit_limits); -1 = no limit"""
    self.lib.b2rt_fg_deposit_limits.argtypes = [C.c_void_p, C.c_int64, C.c_int64]
    self._ck(self.lib.b2rt_fg_deposit_limits(self.h, int(caustic_paths), int(indirect_paths)))


def _irradiance(self, which, queries, k, max_dist2):
    """pbrt EPhoton at the queries' (p, n): (nq, 3) float32."""
    queries = np.ascontiguousarray(queries, QUERY_DTYPE)
    E = np.zeros((len(queries), 3), np.float32)
    self.lib.b2rt_irradiance.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_int32, C.c_float, C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_irradiance(self.h, int(which), len(queries), _ptr(queries), int(k), float(max_dist2), _ptr(E), HOST))
    return E


def _radiance_compute(self, nused=50, max_dist=0.1):
    self.lib.b2rt_radiance_compute.argtypes = [C.c_void_p, C.c_int32, C.c_float]
    self._ck(self.lib.b2rt_radiance_compute(self.h, int(nused), float(np.float32(max_dist) * np.float32(max_dist))))


def _radiance_lookup(self, queries):
    """Nearest radiance photon facing each query's normal: (Lo (nq, 3), index into the sorted radiance photons or -1)."""
    queries = np.ascontiguousarray(queries, QUERY_DTYPE)
    rgb, idx = np.zeros((len(queries), 3), np.float32), np.zeros(len(queries), np.int32)
    self.lib.b2rt_radiance_lookup.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_radiance_lookup(self.h, len(queries), _ptr(queries), _ptr(rgb), _ptr(idx), HOST))
    return rgb, idx


def _set_final_gather(self, enable=True, gather_samples=32, gather_angle=10.0):
    """pbrt `"bool finalgather"`, `"integer finalgathersamples"`, `"float gatherangle"` of the photonmap integrator."""
    self.lib.b2rt_final_gather_set.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_float]
    self._ck(self.lib.b2rt_final_gather_set(self.h, 1 if enable else 0, int(gather_samples), float(gather_angle)))


def _final_gather(self, queries, max_dist2):
    """The camera pass's final gather at caller-supplied shading points (pad = ray eps, pixel id bits, sample word bits)."""
    queries = np.ascontiguousarray(queries, QUERY_DTYPE)
    rgb = np.zeros((len(queries), 3), np.float32)
    self.lib.b2rt_final_gather.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_float, C.c_void_p, C.c_int]
    self._ck(self.lib.b2rt_final_gather(self.h, len(queries), _ptr(queries), float(max_dist2), _ptr(rgb), HOST))
    return rgb


Context.final_gather = _final_gather
Context.photons_trace_fg = _photons_trace_fg
Context.fg_deposit_limits = _fg_deposit_limits
Context.irradiance = _irradiance
Context.radiance_compute = _radiance_compute
Context.radiance_lookup = _radiance_lookup
Context.set_final_gather = _set_final_gather


# ---- several GPUs of one box in ONE process (csrc/multi.cu, include/b2rt.h "b2rt_multi_*") ------------------------------
SYMBOLS += ["b2rt_multi_create", "b2rt_multi_resize", "b2rt_multi_destroy", "b2rt_multi_last_error", "b2rt_multi_count", "b2rt_multi_ctx",
            "b2rt_multi_info", "b2rt_multi_stats_get", "b2rt_multi_build_accel", "b2rt_multi_photons_clear",
            "b2rt_multi_photons_trace", "b2rt_multi_photons_shoot", "b2rt_multi_photons_trace_fg",
            "b2rt_multi_photon_grid_build", "b2rt_multi_radiance_compute", "b2rt_multi_final_gather_set",
            "b2rt_multi_sampler_set", "b2rt_multi_film_set_filter", "b2rt_multi_render_frame"]
EXCHANGE_MODES = {0: "fused peer stores", 1: "nccl broadcast", 2: "peer copies", 3: "push kernel"}


class MultiStats(C.Structure):
    _fields_ = [("trace_ms", C.c_float), ("reserve_ms", C.c_float), ("exchange_ms", C.c_float), ("grid_ms", C.c_float),
                ("render_ms", C.c_float), ("exchange_bytes", C.c_int64), ("passes", C.c_int32)]


class _BorrowedContext(Context):
    """A per-device context owned by a MultiContext: same methods, never destroyed from here."""

    def __init__(self, lib, handle):
        self.lib, self.h = lib, C.c_void_p(handle)

    def close(self):
        self.h = None

    __del__ = close


class MultiContext:
    """n GPUs behind one object.  `ctxs` are the per-device contexts (scene ingestion: the same add_* calls on each, or
    fill(lambda ctx: ...)); the photon passes, grids and frames below are sharded by the library."""

    def __init__(self, ngpus, seed=0, devices=None):
        self.lib = load()
        L, vp, i32, i64, f = self.lib, C.c_void_p, C.c_int32, C.c_int64, C.c_float
        L.b2rt_multi_create.argtypes = [C.c_int, vp, C.c_uint64, C.POINTER(vp)]
        L.b2rt_multi_destroy.argtypes = [vp]
        L.b2rt_multi_destroy.restype = None
        L.b2rt_multi_last_error.argtypes = [vp]
        L.b2rt_multi_last_error.restype = C.c_char_p
        L.b2rt_multi_count.argtypes = [vp]
        L.b2rt_multi_ctx.argtypes = [vp, C.c_int]
        L.b2rt_multi_ctx.restype = vp
        L.b2rt_multi_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32)]
        L.b2rt_multi_stats_get.argtypes = [vp, C.POINTER(MultiStats), C.c_int]
        L.b2rt_multi_build_accel.argtypes = [vp]
        L.b2rt_multi_photons_clear.argtypes = [vp]
        L.b2rt_multi_photons_trace.argtypes = [vp, i64, i64, i32, i32, vp]
        L.b2rt_multi_photons_shoot.argtypes = [vp, i64, i64, i32, i64, vp, vp]
        L.b2rt_multi_photons_trace_fg.argtypes = [vp, i64, i64, i32, vp]
        L.b2rt_multi_photon_grid_build.argtypes = [vp, i32, f, f]
        L.b2rt_multi_radiance_compute.argtypes = [vp, i32, f]
        L.b2rt_multi_final_gather_set.argtypes = [vp, i32, i32, f]
        L.b2rt_multi_sampler_set.argtypes = [vp, i32, i32, i32]
        L.b2rt_multi_film_set_filter.argtypes = [vp, vp]
        L.b2rt_multi_render_frame.argtypes = [vp, C.POINTER(Camera), C.POINTER(FrameParams), vp, C.c_int]
        dev = None if devices is None else (C.c_int * ngpus)(*devices)
        h = vp()
        rc = L.b2rt_multi_create(ngpus, dev, seed, C.byref(h))
        if rc:
            raise B2rtError(rc, L.b2rt_last_error(None).decode())
        self.h = h
        self.ctxs = [_BorrowedContext(L, L.b2rt_multi_ctx(h, i)) for i in range(ngpus)]

    def close(self):
        if getattr(self, "h", None):
            for c in self.ctxs:
                c.close()
            self.lib.b2rt_multi_destroy(self.h)
            self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc):
        if rc:
            raise B2rtError(rc, self.lib.b2rt_multi_last_error(self.h).decode())

    def resize(self, ngpus, devices=None):
        """grow the group; the new contexts copy context 0's scene and settings (then build_accel again)"""
        self.lib.b2rt_multi_resize.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        dev = None if devices is None else (C.c_int * ngpus)(*devices)
        self._ck(self.lib.b2rt_multi_resize(self.h, ngpus, dev))
        self.ctxs = [_BorrowedContext(self.lib, self.lib.b2rt_multi_ctx(self.h, i)) for i in range(ngpus)]

    def fill(self, fn):
        """fn(ctx) on every per-device context (scene ingestion); returns the list of results"""
        return [fn(c) for c in self.ctxs]

    def info(self):
        p, e = C.c_int32(), C.c_int32()
        self._ck(self.lib.b2rt_multi_info(self.h, C.byref(p), C.byref(e)))
        return {"gpus": len(self.ctxs), "peer_access": bool(p.value), "exchange": EXCHANGE_MODES.get(e.value, str(e.value))}

    def stats(self, reset=False):
        s = MultiStats()
        self._ck(self.lib.b2rt_multi_stats_get(self.h, C.byref(s), 1 if reset else 0))
        return {k: getattr(s, k) for k, _ in MultiStats._fields_}

    def build_accel(self):
        self._ck(self.lib.b2rt_multi_build_accel(self.h))
        return self.ctxs[0].accel_info()

    def photons_clear(self):
        self._ck(self.lib.b2rt_multi_photons_clear(self.h))

    def photons_trace(self, first_path, n_paths, max_depth=5, caustic=True, indirect=True):
        st = (C.c_int64 * 2)()
        self._ck(self.lib.b2rt_multi_photons_trace(self.h, first_path, n_paths, max_depth,
                                                   (1 if caustic else 0) | (2 if indirect else 0), st))
        return st[0], st[1]

    def photons_shoot(self, n_caustic, n_indirect, max_depth=5, batch_paths=65536):
        npaths, nst = (C.c_int64 * 2)(), (C.c_int64 * 2)()
        self._ck(self.lib.b2rt_multi_photons_shoot(self.h, n_caustic, n_indirect, max_depth, batch_paths, npaths, nst))
        return (npaths[0], npaths[1]), (nst[0], nst[1])

    def photons_trace_fg(self, first_path, n_paths, max_depth=5):
        st = (C.c_int64 * 2)()
        self._ck(self.lib.b2rt_multi_photons_trace_fg(self.h, int(first_path), int(n_paths), int(max_depth), st))
        return st[0], st[1]

    def grid_build(self, which, cell_size=0.0, target_per_cell=0.0):
        self._ck(self.lib.b2rt_multi_photon_grid_build(self.h, which, cell_size, target_per_cell))

    def radiance_compute(self, nused=50, max_dist=0.1):
        self._ck(self.lib.b2rt_multi_radiance_compute(self.h, int(nused), float(np.float32(max_dist) * np.float32(max_dist))))

    def set_final_gather(self, enable=True, gather_samples=32, gather_angle=10.0):
        self._ck(self.lib.b2rt_multi_final_gather_set(self.h, 1 if enable else 0, int(gather_samples), float(gather_angle)))

    def set_sampler(self, name="random", xsamples=1, ysamples=1):
        self._ck(self.lib.b2rt_multi_sampler_set(self.h, SAMPLERS[name] if isinstance(name, str) else int(name), int(xsamples),
                                                 int(ysamples)))

    def set_filter(self, flt=None):
        self._ck(self.lib.b2rt_multi_film_set_filter(self.h, C.byref(flt) if flt is not None else None))

    def render_frame(self, cam, spp=1, max_specular_depth=5, nused=50, max_dist=0.1, caustic=True, indirect=True,
                     direct=True, window=None, out=None):
        x0, x1, y0, y1 = window if window else (0, cam.xres, 0, cam.yres)
        fp = FrameParams(spp, max_specular_depth, nused, float(np.float32(max_dist) * np.float32(max_dist)), int(caustic),
                         int(indirect), int(direct), x0, x1, y0, y1)
        if out is None:
            out = np.zeros((y1 - y0, x1 - x0, 3), np.float32)
        self._ck(self.lib.b2rt_multi_render_frame(self.h, C.byref(cam), C.byref(fp), _ptr(out), HOST))
        return out
