"""Procedural synthetic inputs: everything is generated from a seed, no assets.

Cfg2 = 1 M-triangle fbm heightfield + primary / shadow / incoherent ray sets.
Cfg4 = same generator at 2000x2000 (8 M triangles).
"""
import numpy as np


def _hash2(ix, iz, seed):
    """32-bit integer hash of a lattice point -> float in [0,1)."""
    h = (ix.astype(np.uint64) * np.uint64(0x9E3779B1) + iz.astype(np.uint64) * np.uint64(0x85EBCA77)
         + np.uint64(seed) * np.uint64(0xC2B2AE3D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & np.uint64(0xFFFFFFFF)
    h ^= h >> np.uint64(15)
    return (h & np.uint64(0xFFFFFF)).astype(np.float64) / float(1 << 24)


def _value_noise(x, z, seed):
    x0 = np.floor(x).astype(np.int64)
    z0 = np.floor(z).astype(np.int64)
    fx, fz = x - x0, z - z0
    sx, sz = fx * fx * (3 - 2 * fx), fz * fz * (3 - 2 * fz)
    a = _hash2(x0, z0, seed)
    b = _hash2(x0 + 1, z0, seed)
    c = _hash2(x0, z0 + 1, seed)
    d = _hash2(x0 + 1, z0 + 1, seed)
    return (a * (1 - sx) + b * sx) * (1 - sz) + (c * (1 - sx) + d * sx) * sz


def fbm(x, z, seed=1234, octaves=4):
    v = np.zeros_like(x, dtype=np.float64)
    amp, freq = 0.5, 4.0
    for o in range(octaves):
        v += amp * (_value_noise(x * freq, z * freq, seed + o) * 2 - 1)
        amp *= 0.5
        freq *= 2.0
    return v


def heightfield_mesh(nx, nz, ntris=None, seed=1234, shuffle=True):
    """(nx+1)x(nz+1) vertex grid over [-1,1]^2, y = 0.15*fbm; 2*nx*nz triangles (+ a 2-triangle ground quad
    at the end), trimmed to exactly `ntris`, triangle order shuffled (seeded) so it is not spatially sorted."""
    xs = np.linspace(-1.0, 1.0, nx + 1)
    zs = np.linspace(-1.0, 1.0, nz + 1)
    X, Z = np.meshgrid(xs, zs, indexing="xy")          # (nz+1, nx+1)
    Y = 0.15 * fbm(X, Z, seed)
    P = np.stack([X, Y, Z], -1).reshape(-1, 3).astype(np.float32)
    i = np.arange(nx)[None, :] + (nx + 1) * np.arange(nz)[:, None]
    a, b, c, d = i, i + 1, i + (nx + 1), i + (nx + 2)
    tris = np.stack([np.stack([a, c, b], -1), np.stack([b, c, d], -1)], 2).reshape(-1, 3)
    # ground quad below the terrain
    g0 = len(P)
    ground = np.array([[-1.5, -0.4, -1.5], [1.5, -0.4, -1.5], [1.5, -0.4, 1.5], [-1.5, -0.4, 1.5]], np.float32)
    P = np.concatenate([P, ground])
    gt = np.array([[g0, g0 + 2, g0 + 1], [g0, g0 + 3, g0 + 2]])
    if ntris is not None:
        tris = tris[: ntris - 2]
    tris = np.concatenate([tris, gt]).astype(np.int32)
    if shuffle:
        rng = np.random.default_rng(seed)
        tris = tris[rng.permutation(len(tris))]
    return P, np.ascontiguousarray(tris)


def cfg2_mesh(ntris=1_000_000, seed=1234):
    """708x707 grid trimmed to exactly 1 000 000 triangles."""
    if ntris == 1_000_000:
        return heightfield_mesh(708, 707, ntris, seed)
    side = int(np.ceil(np.sqrt(ntris / 2)))
    return heightfield_mesh(side, side, ntris, seed)


def cfg4_mesh(seed=1234):
    return heightfield_mesh(2000, 2000, 8_000_000, seed)


def pack_rays(o, d, tmin, tmax):
    n = len(o)
    r = np.empty((n, 8), np.float32)
    r[:, 0:3] = o
    r[:, 3] = tmin
    r[:, 4:7] = d
    r[:, 7] = tmax
    return r


def pinhole_rays(width, height, eye=(0.0, 1.2, 2.5), look=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0), fov=45.0,
                 spp=1, seed=7):
    """Primary rays through pixel centres (spp=1) or seeded jitter (spp>1); normalised directions."""
    eye = np.asarray(eye, np.float64)
    f = np.asarray(look, np.float64) - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, np.asarray(up, np.float64))
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    th = np.tan(np.radians(fov) / 2)
    aspect = width / height
    py, px = np.meshgrid(np.arange(height), np.arange(width), indexing="ij")
    px = np.repeat(px.reshape(-1), spp).astype(np.float64)
    py = np.repeat(py.reshape(-1), spp).astype(np.float64)
    if spp == 1:
        jx = jy = 0.5
    else:
        rng = np.random.default_rng(seed)
        jx, jy = rng.random(len(px)), rng.random(len(px))
    sx = (2 * (px + jx) / width - 1) * th * aspect
    sy = (1 - 2 * (py + jy) / height) * th
    d = f[None, :] + sx[:, None] * s[None, :] + sy[:, None] * u[None, :]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    o = np.broadcast_to(eye, d.shape)
    return pack_rays(o, d, 0.0, np.inf)


def incoherent_rays(n, lo, hi, seed=99):
    """Origins uniform in the box, directions uniform on the sphere (aggregatetest-style,
    pbrt-v2/src/renderers/aggregatetest.cpp:74-83)."""
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    o = lo + rng.random((n, 3)) * (hi - lo)
    z = 1 - 2 * rng.random(n)
    r = np.sqrt(np.maximum(0, 1 - z * z))
    phi = 2 * np.pi * rng.random(n)
    d = np.stack([r * np.cos(phi), r * np.sin(phi), z], -1)
    return pack_rays(o, d, 0.0, np.inf)


def shadow_rays(primary, hits_t, hits_prim, light=(0.5, 3.0, 0.5)):
    """Segment rays from each primary hit to a point light: d = light - p (unnormalised), tmin 1e-3*t... the
    reference's form (cuda_render/simple_render/simplerender.cu:65): tmin 1e-3, tmax 1-1e-3 on the segment."""
    ok = hits_prim >= 0
    o = primary[ok, 0:3].astype(np.float64) + primary[ok, 4:7].astype(np.float64) * hits_t[ok, None].astype(np.float64)
    d = np.asarray(light, np.float64)[None, :] - o
    return pack_rays(o, d, 1e-3, 1.0 - 1e-3)


def cornell_geometry():
    """Appendix D Cornell box as flat arrays: list of (P, idx, rgb) quads + two boxes."""
    def quad(p):
        return np.asarray(p, np.float32), np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    white, red, green = (0.73, 0.73, 0.73), (0.65, 0.05, 0.05), (0.12, 0.45, 0.15)
    parts = [
        (*quad([[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1]]), white),      # floor
        (*quad([[-1, 2, -1], [-1, 2, 1], [1, 2, 1], [1, 2, -1]]), white),      # ceiling
        (*quad([[-1, 0, -1], [-1, 2, -1], [1, 2, -1], [1, 0, -1]]), white),    # back
        (*quad([[-1, 0, -1], [-1, 0, 1], [-1, 2, 1], [-1, 2, -1]]), red),      # left
        (*quad([[1, 0, -1], [1, 2, -1], [1, 2, 1], [1, 0, 1]]), green),        # right
    ]
    return parts


# ---- full scene descriptions (geometry + materials + lights), neutral format shared by the b2rt binding, the C oracle
# ---- and the .pbrt writer ------------------------------------------------------------------------------------------
MATTE, MIRROR, GLASS = 0, 1, 2


def xf_translate(tx, ty, tz):
    o2w = np.array([1, 0, 0, tx, 0, 1, 0, ty, 0, 0, 1, tz], np.float32)
    w2o = np.array([1, 0, 0, -tx, 0, 1, 0, -ty, 0, 0, 1, -tz], np.float32)
    return o2w, w2o


def xf_disk_down(x, y, z):
    """pbrt `Translate x y z  Rotate 90 1 0 0`: object z axis -> world -y (a ceiling light facing down)."""
    o2w = np.array([1, 0, 0, x, 0, 0, -1, y, 0, 1, 0, z], np.float32)
    w2o = np.array([1, 0, 0, -x, 0, 0, 1, -z, 0, -1, 0, y], np.float32)
    return o2w, w2o


def _box(lo, hi):
    """axis-aligned box as 12 outward-facing triangles"""
    x0, y0, z0 = lo
    x1, y1, z1 = hi
    P = np.array([[x0, y0, z0], [x1, y0, z0], [x1, y1, z0], [x0, y1, z0],
                  [x0, y0, z1], [x1, y0, z1], [x1, y1, z1], [x0, y1, z1]], np.float32)
    q = [[0, 3, 2, 1], [4, 5, 6, 7], [0, 1, 5, 4], [3, 7, 6, 2], [0, 4, 7, 3], [1, 2, 6, 5]]
    idx = np.array([[a, b, c] for a, b, c, d in q] + [[a, c, d] for a, b, c, d in q], np.int32)
    return P, idx


def cornell_scene(spheres=True, boxes=False, wide=1.0, light="disk"):
    """SURVEY Appendix D.  spheres: glass + mirror sphere (Cfg3 flavour); boxes: the two matte boxes (Cfg1 flavour);
    wide: x half-extent of the box (1.78 for the 16:9 Cfg3 frame)."""
    W = float(wide)
    mats = [(MATTE, (0.73, 0.73, 0.73), (0, 0, 0), 1.5), (MATTE, (0.65, 0.05, 0.05), (0, 0, 0), 1.5),
            (MATTE, (0.12, 0.45, 0.15), (0, 0, 0), 1.5), (GLASS, (1, 1, 1), (1, 1, 1), 1.5),
            (MIRROR, (0.9, 0.9, 0.9), (0, 0, 0), 1.5)]
    q = np.array([[0, 1, 2], [0, 2, 3]], np.int32)

    def quad(p):
        return np.asarray(p, np.float32)
    shapes = [
        ("mesh", quad([[-W, 0, -1], [W, 0, -1], [W, 0, 1], [-W, 0, 1]]), q, 0),       # floor
        ("mesh", quad([[-W, 2, -1], [-W, 2, 1], [W, 2, 1], [W, 2, -1]]), q, 0),       # ceiling
        ("mesh", quad([[-W, 0, -1], [-W, 2, -1], [W, 2, -1], [W, 0, -1]]), q, 0),     # back
        ("mesh", quad([[-W, 0, -1], [-W, 0, 1], [-W, 2, 1], [-W, 2, -1]]), q, 1),     # left (red)
        ("mesh", quad([[W, 0, -1], [W, 2, -1], [W, 2, 1], [W, 0, 1]]), q, 2),         # right (green)
    ]
    if boxes:
        shapes.append(("mesh", *_box((-0.7, 0.0, -0.6), (-0.15, 1.2, -0.05)), 0))
        shapes.append(("mesh", *_box((0.15, 0.0, 0.0), (0.7, 0.6, 0.55)), 0))
    if spheres:
        shapes.append(("sphere", *xf_translate(0.4, 0.35, 0.3), 0.35, 3))
        shapes.append(("sphere", *xf_translate(-0.45, 0.35, -0.3), 0.35, 4))
    lights = []
    if light == "disk":
        shapes.append(("disk", *xf_disk_down(0.0, 1.98, 0.0), 0.0, 0.3, 0.0, 0))
        lights.append(("disk", len(shapes) - 1, (17.0, 12.0, 4.0), 4))
    elif light == "spot":           # pbrt `LightSource "spot"` under the ceiling, looking down at the floor
        lights.append(spot_light((0.2, 1.9, 0.1), (0.0, 0.0, 0.0), 50.0, 10.0, (9.0, 8.0, 7.0)))
    elif light == "distant":        # pbrt `LightSource "distant"` shining in through the open front of the box
        lights.append(distant_light((0.4, 0.9, 3.0), (0.0, 0.6, 0.0), (2.5, 2.5, 2.0)))
    elif light == "mixed":          # one of each delta light
        lights.append(("point", (-0.5, 1.7, 0.3), (1.5, 1.5, 1.5)))
        lights.append(spot_light((0.2, 1.9, 0.1), (0.0, 0.0, 0.0), 50.0, 10.0, (9.0, 8.0, 7.0)))
        lights.append(distant_light((0.4, 0.9, 3.0), (0.0, 0.6, 0.0), (2.5, 2.5, 2.0)))
    else:
        lights.append(("point", (0.0, 1.9, 0.0), (3.0, 3.0, 3.0)))
    return {"materials": mats, "shapes": shapes, "lights": lights}


def _coordinate_system(v1):
    """pbrt CoordinateSystem (core/geometry.h:508-518) in float32"""
    v1 = np.asarray(v1, np.float32)
    if abs(v1[0]) > abs(v1[1]):
        inv = np.float32(1) / np.sqrt(v1[0] * v1[0] + v1[2] * v1[2], dtype=np.float32)
        v2 = np.array([-v1[2] * inv, 0, v1[0] * inv], np.float32)
    else:
        inv = np.float32(1) / np.sqrt(v1[1] * v1[1] + v1[2] * v1[2], dtype=np.float32)
        v2 = np.array([0, v1[2] * inv, -v1[1] * inv], np.float32)
    return v2, np.cross(v1.astype(np.float64), v2.astype(np.float64)).astype(np.float32)


def spot_light(frm, to, coneangle, conedelta, rgb):
    """("spot", l2w, w2l, rgb, cos_total_width, cos_falloff_start, from, to, coneangle, conedelta): pbrt CreateSpotLight
    (lights/spot.cpp:78-101) with an identity CTM -- LightToWorld = Translate(from) * Inverse(dirToZ)"""
    frm, to = np.asarray(frm, np.float32), np.asarray(to, np.float32)
    d = (to - frm).astype(np.float32)
    d = (d / np.sqrt((d * d).sum(dtype=np.float32), dtype=np.float32)).astype(np.float32)
    du, dv = _coordinate_system(d)
    R = np.stack([du, dv, d], 1).astype(np.float32)                 # columns: light x, y, z axes in world space
    l2w = np.concatenate([R, frm[:, None]], 1).astype(np.float32)
    w2l = np.concatenate([R.T, (-R.T.astype(np.float64) @ frm.astype(np.float64))[:, None].astype(np.float32)], 1).astype(np.float32)
    ct = np.float32(np.cos(np.float32(np.radians(np.float32(coneangle)))))
    cf = np.float32(np.cos(np.float32(np.radians(np.float32(coneangle - conedelta)))))
    return ("spot", l2w.reshape(-1), w2l.reshape(-1), tuple(rgb), float(ct), float(cf), tuple(frm), tuple(to), coneangle, conedelta)


def distant_light(frm, to, rgb):
    """("distant", dir_to_light, rgb, from, to): pbrt CreateDistantLight (lights/distant.cpp:65-73), lightDir = Normalize(from - to)"""
    d = (np.asarray(frm, np.float32) - np.asarray(to, np.float32)).astype(np.float32)
    d = (d * (np.float32(1) / np.sqrt((d * d).sum(dtype=np.float32), dtype=np.float32))).astype(np.float32)
    return ("distant", tuple(float(x) for x in d), tuple(rgb), tuple(frm), tuple(to))


def fill_scene(target, desc):
    """target: anything with add_material / add_mesh / add_sphere / add_disk / add_light in the b2rt.Context signature
    (b2rt.Context itself, or oracle.PortScene).  Returns the first prim id of every shape."""
    mids = [target.add_material(t, kd, kt, ior) for (t, kd, kt, ior) in desc["materials"]]
    first = []
    light_of_shape = {l[1]: i for i, l in enumerate(desc["lights"]) if l[0] == "disk"}
    for si, sh in enumerate(desc["shapes"]):
        al = light_of_shape.get(si, -1)
        if sh[0] == "mesh":
            if len(sh) > 4:       # per-vertex normals / uv (world space here: descriptions with an o2w are for to_pbrt only)
                assert len(sh) < 7 or sh[6] is None
                first.append(target.add_mesh(sh[1], sh[2], material=mids[sh[3]], arealight=al, N=sh[4], uv=sh[5] if len(sh) > 5 else None))
            else:
                first.append(target.add_mesh(sh[1], sh[2], material=mids[sh[3]], arealight=al))
        elif sh[0] == "sphere":
            first.append(target.add_sphere(sh[1], sh[3], w2o=sh[2], material=mids[sh[4]], arealight=al))
        else:
            first.append(target.add_disk(sh[1], sh[3], sh[4], sh[5], w2o=sh[2], material=mids[sh[6]], arealight=al))
    for l in desc["lights"]:
        if l[0] == "disk":
            target.add_light(1, l[2], prim=first[l[1]], nsamples=l[3])
        elif l[0] == "spot":
            target.add_spot_light(l[1], l[2], l[3], l[4], l[5])
        elif l[0] == "distant":
            target.add_distant_light(l[1], l[2])
        else:
            target.add_light(0, l[2], p=l[1])
    return first


# ---- .pbrt text writer (for the reference's CPU renderer and for the drop-in run) and PFM reader ------------------------
def _m16(o2w):
    """row-major 3x4 -> pbrt `Transform [16]` (column-major 4x4)"""
    m = np.vstack([np.asarray(o2w, np.float64).reshape(3, 4), [0, 0, 0, 1]])
    return " ".join("%.9g" % v for v in m.T.reshape(-1))


def to_pbrt(desc, eye, look, up, fov, xres, yres, spp, integrator, filename, renderer=None, sampler="lowdiscrepancy",
            pixelfilter='"box"', world_rotate_y=0.0, extra_world=""):
    """integrator: the full `SurfaceIntegrator ...` line.  Only features the reference's GPU path translates are used.
    extra_world: raw text appended inside the world block.
    world_rotate_y (degrees): the whole world AND the camera are turned about the y axis -- the same picture in expectation,
    but pbrt's constant-seeded random streams (photon directions, light samples) then produce different paths: an
    independent realisation of the reference renderer (tools/converged_parity.py averages several)."""
    rot = ""
    if world_rotate_y:
        a = np.deg2rad(world_rotate_y)
        R = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]])      # pbrt Rotate(theta, y)
        eye, look, up = R @ np.asarray(eye, float), R @ np.asarray(look, float), R @ np.asarray(up, float)
        rot = "Rotate %.9g 0 1 0" % world_rotate_y
    xf = "ConcatTransform" if world_rotate_y else "Transform"
    L = ["LookAt %g %g %g  %g %g %g  %g %g %g" % (*eye, *look, *up),
         'Camera "perspective" "float fov" [%g]' % fov,
         'Film "image" "integer xresolution" [%d] "integer yresolution" [%d] "string filename" ["%s"]' % (xres, yres, filename),
         ('Sampler "stratified" "integer xsamples" [%d] "integer ysamples" [%d]' % (int(round(spp ** 0.5)), int(round(spp ** 0.5)))
          if sampler == "stratified" else 'Sampler "%s" "integer pixelsamples" [%d]' % (sampler, spp)),
         'PixelFilter ' + pixelfilter, integrator]
    if renderer:
        L.append(renderer if renderer.startswith("Renderer") else 'Renderer "%s"' % renderer)
    L.append("WorldBegin")
    if rot:
        L.append(rot)
    light_of_shape = {l[1]: l for l in desc["lights"] if l[0] == "disk"}
    for l in desc["lights"]:
        if l[0] == "point":
            L.append('LightSource "point" "point from" [%g %g %g] "rgb I" [%g %g %g]' % (*l[1], *l[2]))
        elif l[0] == "spot":
            L.append('LightSource "spot" "point from" [%.9g %.9g %.9g] "point to" [%.9g %.9g %.9g] "float coneangle" [%g] '
                     '"float conedeltaangle" [%g] "rgb I" [%g %g %g]' % (*l[6], *l[7], l[8], l[9], *l[3]))
        elif l[0] == "distant":
            L.append('LightSource "distant" "point from" [%.9g %.9g %.9g] "point to" [%.9g %.9g %.9g] "rgb L" [%g %g %g]'
                     % (*l[3], *l[4], *l[2]))
    for si, sh in enumerate(desc["shapes"]):
        mid = sh[3] if sh[0] == "mesh" else (sh[4] if sh[0] == "sphere" else sh[6])
        t, kd, kt, ior = desc["materials"][mid]
        L.append("AttributeBegin")
        if t == MATTE:
            L.append('  Material "matte" "rgb Kd" [%g %g %g]' % tuple(kd))
        elif t == MIRROR:
            L.append('  Material "mirror" "rgb Kr" [%g %g %g]' % tuple(kd))
        else:
            L.append('  Material "glass" "rgb Kr" [%g %g %g] "rgb Kt" [%g %g %g] "float index" [%g]' % (*kd, *kt, ior))
        if si in light_of_shape:
            l = light_of_shape[si]
            L.append('  AreaLightSource "diffuse" "rgb L" [%g %g %g] "integer nsamples" [%d]' % (*l[2], l[3]))
        if sh[0] == "mesh":
            # optional per-vertex attributes: ("mesh", P, idx, material, N, uv[, o2w]) -- with o2w, P and N are in object
            # space and the file carries the transform (pbrt transforms P at load time and N per hit)
            extra = ""
            if len(sh) > 4 and sh[4] is not None:
                extra += ' "normal N" [%s]' % " ".join("%.9g" % v for v in np.asarray(sh[4], np.float64).reshape(-1))
            if len(sh) > 5 and sh[5] is not None:
                extra += ' "float uv" [%s]' % " ".join("%.9g" % v for v in np.asarray(sh[5], np.float64).reshape(-1))
            if len(sh) > 6 and sh[6] is not None:
                L.append("  %s [%s]" % (xf, _m16(sh[6])))
            L.append('  Shape "trianglemesh" "integer indices" [%s] "point P" [%s]%s' % (
                " ".join(str(int(v)) for v in np.asarray(sh[2]).reshape(-1)),
                " ".join("%.9g" % v for v in np.asarray(sh[1], np.float64).reshape(-1)), extra))
        elif sh[0] == "sphere":
            L.append('  %s [%s]\n  Shape "sphere" "float radius" [%g]' % (xf, _m16(sh[1]), sh[3]))
        else:
            L.append('  %s [%s]\n  Shape "disk" "float height" [%g] "float radius" [%g] "float innerradius" [%g]' % (
                xf, _m16(sh[1]), sh[3], sh[4], sh[5]))
        L.append("AttributeEnd")
    if extra_world:                 # raw scene text (ObjectBegin / ObjectInstance blocks of the instancing tests)
        L.append(extra_world)
    L.append("WorldEnd")
    return "\n".join(L) + "\n"


def read_pfm(path):
    with open(path, "rb") as f:
        kind = f.readline().strip()
        w, h = map(int, f.readline().split())
        scale = float(f.readline())
        data = np.frombuffer(f.read(), "<f4" if scale < 0 else ">f4")
    img = data.reshape(h, w, 3 if kind == b"PF" else 1)
    return np.ascontiguousarray(img[::-1]).astype(np.float32)      # PFM stores bottom row first


def image_metrics(img, ref):
    """exrdiff-style metrics (pbrt-v2/src/tools/exrdiff.cpp:73-103), normalised by the reference's mean:
    mean |a-b| / mean b  (mean relative error),  RMSE / mean b,  and the bias of the image mean."""
    a, b = np.asarray(img, np.float64), np.asarray(ref, np.float64)
    mb = b.mean()
    return {"mean_rel": float(np.abs(a - b).mean() / mb), "rmse_rel": float(np.sqrt(((a - b) ** 2).mean()) / mb),
            "bias_rel": float((a.mean() - mb) / mb)}


def sphere_mesh(nu=24, nv=12, radius=0.5, center=(0.0, 0.0, 0.0), seed=None):
    """Latitude / longitude sphere as an indexed mesh with per-vertex normals (exact sphere normals: the mesh shades
    smooth) and uv = (phi / 2 pi, theta / pi) -- pbrt's "N" and "uv" mesh parameters.  The poles are rings of
    coincident vertices so that every triangle has three distinct uv.  seed: shuffle the triangle order."""
    th = np.linspace(0.0, np.pi, nv + 1)
    ph = np.linspace(0.0, 2.0 * np.pi, nu + 1)
    T, Ph = np.meshgrid(th, ph, indexing="ij")
    n = np.stack([np.sin(T) * np.cos(Ph), np.cos(T), np.sin(T) * np.sin(Ph)], -1).reshape(-1, 3)
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    P = (np.asarray(center, np.float64) + radius * n).astype(np.float32)
    uv = np.stack([Ph / (2.0 * np.pi), T / np.pi], -1).reshape(-1, 2).astype(np.float32)
    idx = []
    for i in range(nv):
        for j in range(nu):
            a, b = i * (nu + 1) + j, i * (nu + 1) + j + 1
            c, d = a + nu + 1, b + nu + 1
            if i > 0:
                idx.append((a, b, c))
            if i < nv - 1:
                idx.append((b, d, c))
    idx = np.asarray(idx, np.int32)
    if seed is not None:
        idx = idx[np.random.default_rng(seed).permutation(len(idx))]
    return P, idx, n.astype(np.float32), uv


# ---- object instances (two-level BVH) ----------------------------------------------------------------------------------
def xf_compose(*ms):
    """product of 3x4 row-major affine transforms (applied right to left), float64 -> float32, with its inverse"""
    M = np.eye(4)
    for m in ms:
        M = M @ np.vstack([np.asarray(m, np.float64).reshape(3, 4), [0, 0, 0, 1]])
    return M[:3].astype(np.float32).reshape(12), np.linalg.inv(M)[:3].astype(np.float32).reshape(12)


def xf_rotate_scale(axis, deg, scale, t):
    """T(t) * R(axis, deg) * S(scale): (o2w, w2o) as float32[12]"""
    a = np.asarray(axis, np.float64); a /= np.linalg.norm(a)
    c, s = np.cos(np.radians(deg)), np.sin(np.radians(deg))
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    R = np.eye(3) * c + s * K + (1 - c) * np.outer(a, a)
    M = np.eye(4); M[:3, :3] = R @ np.diag(np.broadcast_to(np.asarray(scale, np.float64), 3)); M[:3, 3] = t
    return M[:3].astype(np.float32).reshape(12), np.linalg.inv(M)[:3].astype(np.float32).reshape(12)


def instanced_scene(n_inst=10, seed=5, exact=False, attrs=True):
    """World: a ground heightfield and one analytic sphere.  Object 0: a smooth sphere mesh (per-vertex normals + uv when
    attrs) over a flat-shaded heightfield patch.  Object 1: an analytic sphere, a disk and a small mesh.  n_inst instances of
    the two objects under rotations, non-uniform scales and translations.
    exact=True: mesh-only objects with vertices on a 2^-10 grid and integer translations only -- every instance-space
    operation is then exact and the two-level scene returns the same bits as its flattened copy.
    Returns {"world": [shape...], "objects": [[shape...], ...], "instances": [(object, o2w, w2o), ...]} with shapes
    ("mesh", P, idx, N, uv) | ("sphere", o2w, w2o, radius) | ("disk", o2w, w2o, height, radius, inner)."""
    rng = np.random.default_rng(seed)
    q = (lambda a: (np.round(np.asarray(a, np.float64) * 1024.0) / 1024.0).astype(np.float32)) if exact else (lambda a: np.asarray(a, np.float32))
    Pg, ig = heightfield_mesh(16, 16, None, seed=seed + 1)
    Pg = Pg.copy(); Pg[:, 1] = Pg[:, 1] * 0.3 - 0.8; Pg[:, [0, 2]] *= 4.0
    world = [("mesh", q(Pg), ig, None, None)]
    if not exact:
        o2w, w2o = xf_translate(-2.5, 0.2, 1.0)
        world.append(("sphere", o2w, w2o, 0.45))
    P, idx, N, uv = sphere_mesh(16, 8, 0.4, (0.0, 0.3, 0.0), seed=seed + 2)
    Ph, ih = heightfield_mesh(6, 6, None, seed=seed + 3)
    Ph = Ph.copy(); Ph[:, 1] = Ph[:, 1] * 0.2 - 0.2; Ph[:, [0, 2]] *= 0.5
    use = attrs and not exact
    obj0 = [("mesh", q(P), idx, N if use else None, uv if use else None), ("mesh", q(Ph), ih, None, None)]
    objects = [obj0]
    if not exact:
        so2w, sw2o = xf_rotate_scale((1, 2, 0.5), 25.0, (1.0, 1.0, 1.0), (0.1, 0.25, 0.0))
        do2w, dw2o = xf_rotate_scale((1, 0, 0), -90.0, (1.0, 1.0, 1.0), (0.0, -0.1, 0.0))
        Pm, im = heightfield_mesh(4, 4, None, seed=seed + 4)
        Pm = Pm.copy(); Pm[:, 1] = Pm[:, 1] * 0.15 + 0.55; Pm[:, [0, 2]] *= 0.35
        objects.append([("sphere", so2w, sw2o, 0.3), ("disk", do2w, dw2o, 0.0, 0.5, 0.1), ("mesh", Pm.astype(np.float32), im, None, None)])
    inst = []
    for k in range(n_inst):
        o = int(k % len(objects))
        if exact:
            t = rng.integers(-3, 4, 3).astype(np.float64); t[1] = float(rng.integers(0, 3))
            o2w, w2o = xf_translate(*t)
        else:
            t = rng.uniform(-3.0, 3.0, 3); t[1] = rng.uniform(-0.3, 1.2)
            sc = rng.uniform(0.5, 1.6, 3) if k % 3 else np.full(3, rng.uniform(0.6, 1.4))
            if k % 5 == 4:
                sc[0] = -sc[0]                     # a mirroring instance (swaps handedness)
            o2w, w2o = xf_rotate_scale(rng.normal(size=3), rng.uniform(0, 360), sc, t)
        inst.append((o, o2w, w2o))
    return {"world": world, "objects": objects, "instances": inst}


def _add_shape(target, sh, style):
    if sh[0] == "mesh":
        kw = {}
        if sh[3] is not None: kw["N"] = sh[3]
        if sh[4] is not None: kw["uv"] = sh[4]
        if style == "ref":
            return target.add_mesh_attr(sh[1], sh[2], sh[3], sh[4]) if kw else target.add_mesh(sh[1], sh[2])
        return target.add_mesh(sh[1], sh[2], **kw)
    if sh[0] == "sphere":
        return target.add_sphere(sh[1], sh[3], w2o=sh[2]) if style == "b2rt" else target.add_sphere(sh[1], sh[2], sh[3])
    return target.add_disk(sh[1], sh[3], sh[4], sh[5], w2o=sh[2]) if style == "b2rt" else target.add_disk(sh[1], sh[2], sh[3], sh[4], sh[5])


def fill_instanced(target, desc, style="b2rt"):
    """style "b2rt": b2rt.Context signatures (add_sphere(o2w, radius, w2o=)); "port" / "ref": oracle.PortInstancedScene /
    oracle.RefScene (add_sphere(o2w, w2o, radius)).  World shapes first, then the objects, then the instances."""
    for sh in desc["world"]:
        _add_shape(target, sh, style)
    ids = []
    for shapes in desc["objects"]:
        ids.append(target.object_begin())
        for sh in shapes:
            _add_shape(target, sh, style)
        target.object_end()
    return [target.add_instance(ids[o], o2w, w2o) if style != "b2rt" else target.add_instance(ids[o], o2w, w2o=w2o)
            for (o, o2w, w2o) in desc["instances"]]


def flatten_instanced(desc):
    """the same scene with every instance written out as world-space shapes (what the drop-in used to do): world shapes,
    then instance 0's shapes, instance 1's, ...  Points by the float32 expression of pbrt's Transform (exact for the
    `exact` scenes), normals by the inverse transpose, analytic shapes by composing the matrices."""
    def xf_point32(m, x, y, z):          # pbrt Transform(Point) in binary32, one rounding per operation, left to right
        m = np.asarray(m, np.float32).reshape(12)
        return (m[0] * x + m[1] * y + m[2] * z + m[3], m[4] * x + m[5] * y + m[6] * z + m[7], m[8] * x + m[9] * y + m[10] * z + m[11])
    out = list(desc["world"])
    for (o, o2w, w2o) in desc["instances"]:
        for sh in desc["objects"][o]:
            if sh[0] == "mesh":
                P = np.asarray(sh[1], np.float32)
                x, y, z = xf_point32(o2w, P[:, 0], P[:, 1], P[:, 2])
                N = None
                if sh[3] is not None:
                    n = np.asarray(sh[3], np.float32); m = np.asarray(w2o, np.float32).reshape(3, 4)
                    N = np.stack([m[0, 0] * n[:, 0] + m[1, 0] * n[:, 1] + m[2, 0] * n[:, 2], m[0, 1] * n[:, 0] + m[1, 1] * n[:, 1] + m[2, 1] * n[:, 2],
                                  m[0, 2] * n[:, 0] + m[1, 2] * n[:, 1] + m[2, 2] * n[:, 2]], -1).astype(np.float32)
                out.append(("mesh", np.stack([x, y, z], -1).astype(np.float32), sh[2], N, sh[4]))
            else:
                a, b = xf_compose(o2w, sh[1])
                out.append((sh[0], a, b) + tuple(sh[3:]))
    return out
