"""ctypes front-end of the CPU oracle (oracle/rt_oracle.cpp).

TEST INFRASTRUCTURE ONLY -- "parity unpinned" (see the header of rt_oracle.cpp).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "librt_oracle.so")
_lib = None


class OrcParams(C.Structure):
    _fields_ = [
        ("width", C.c_uint32), ("height", C.c_uint32), ("nb_ray", C.c_uint32), ("nb_light_sample", C.c_uint32),
        ("eye", C.c_float * 3), ("look_at", C.c_float * 3), ("up", C.c_float * 3), ("distance", C.c_float),
        ("light", C.c_float * 9),
        ("samples", C.c_void_p), ("n_samples", C.c_uint32),
        ("mode", C.c_int32), ("n_threads", C.c_int32),
    ]


class OrcStats(C.Structure):
    _fields_ = [
        ("primary_rays", C.c_uint64), ("primary_hits", C.c_uint64), ("shadow_rays", C.c_uint64),
        ("shadow_occluded", C.c_uint64), ("box_tests", C.c_uint64), ("tri_tests", C.c_uint64),
        ("nan_asserts", C.c_uint64), ("seconds", C.c_double), ("threads", C.c_int32),
    ]


def build(force=False):
    """Compile librt_oracle.so with the committed Makefile (gcc only, no deps)."""
    src = os.path.join(_HERE, "rt_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "librt_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_create.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_set_materials.argtypes = [C.c_void_p, C.c_void_p]; L.orc_scene_set_materials.restype = None
        L.orc_scene_num_nodes.argtypes = [C.c_void_p]
        L.orc_scene_dfs_rank.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_tri_normal.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_camera_basis.argtypes = [C.c_void_p] * 6
        L.orc_light_sample.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_void_p]
        L.orc_box_intersect.argtypes = [C.c_void_p] * 4
        L.orc_tri_intersect.argtypes = [C.c_void_p] * 4
        L.orc_primary_ray.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_gamma_u8.restype = C.c_uint8
        L.orc_gamma_u8.argtypes = [C.c_float]
        L.orc_sphere_intersect.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_sphere_box.argtypes = [C.c_void_p, C.c_float, C.c_void_p]; L.orc_sphere_box.restype = None
        L.orc_sphere_normal.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]; L.orc_sphere_normal.restype = None
        L.orc_spheres_closest.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_render.restype = C.c_int32
        L.orc_render.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64] + [C.c_void_p] * 6
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


MODE_REFERENCE_TREE = 0
MODE_BRUTE = 1
MODE_BRUTE_ANYHIT = 2


class OracleScene:
    """verts: (n,9) float32 v0|v1|v2, colors: (n,3) float32, in the reference's primitive order."""

    def __init__(self, verts, colors, build_tree=True):
        self.verts = _f32(verts).reshape(-1, 9)
        self.colors = _f32(colors).reshape(-1, 3)
        self.n = self.verts.shape[0]
        self.has_tree = bool(build_tree)
        self._h = lib().orc_scene_create(_ptr(self.verts), _ptr(self.colors), self.n, int(build_tree))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scene_destroy(self._h)
            self._h = None

    def num_nodes(self):
        return lib().orc_scene_num_nodes(self._h)

    def dfs_rank(self):
        r = np.zeros(self.n, dtype=np.int32)
        if lib().orc_scene_dfs_rank(self._h, _ptr(r)) != 0:
            raise RuntimeError("reference tree was not built")
        return r

    def set_materials(self, ks_ns):
        """Phong extension (not in the reference): (n,4) {Ks.rgb, Ns} or None."""
        if ks_ns is None:
            lib().orc_scene_set_materials(self._h, None)
        else:
            k = _f32(ks_ns).reshape(-1, 4)
            assert k.shape[0] == self.n
            lib().orc_scene_set_materials(self._h, _ptr(k))

    def normal(self, p):
        out = np.zeros(3, dtype=np.float32)
        lib().orc_tri_normal(self._h, int(p), _ptr(out))
        return out

    def render(self, width, height, samples, eye, look_at, up, distance, light, nb_ray=1, nb_light_sample=100,
               mode=MODE_REFERENCE_TREE, pixels=None, n_threads=0):
        """Returns dict(rgb_lin, rgb8, hit_count, prim0, t0, stats).  pixels: optional (m,2) uint32 (px,py)."""
        samples = _f32(samples).reshape(-1, 2)
        prm = OrcParams()
        prm.width, prm.height, prm.nb_ray, prm.nb_light_sample = width, height, nb_ray, nb_light_sample
        prm.eye[:] = list(map(float, eye)); prm.look_at[:] = list(map(float, look_at)); prm.up[:] = list(map(float, up))
        prm.distance = float(distance)
        prm.light[:] = [float(x) for x in np.asarray(light, dtype=np.float32).reshape(9)]
        prm.samples = samples.ctypes.data
        prm.n_samples = samples.shape[0]
        prm.mode = mode
        prm.n_threads = n_threads
        if pixels is None:
            m = width * height
            pix_ptr = None
        else:
            pixels = np.ascontiguousarray(pixels, dtype=np.uint32).reshape(-1, 2)
            m = pixels.shape[0]
            pix_ptr = _ptr(pixels)
        out = dict(rgb_lin=np.zeros((m, 3), np.float32), rgb8=np.zeros((m, 3), np.uint8),
                   hit_count=np.zeros(m, np.uint8), prim0=np.zeros(m, np.int32), t0=np.zeros(m, np.float32))
        st = OrcStats()
        rc = lib().orc_render(self._h, C.byref(prm), pix_ptr, m, _ptr(out["rgb_lin"]), _ptr(out["rgb8"]),
                              _ptr(out["hit_count"]), _ptr(out["prim0"]), _ptr(out["t0"]), C.byref(st))
        if rc != 0:
            raise RuntimeError("orc_render failed: %d" % rc)
        out["stats"] = {k: getattr(st, k) for k, _ in OrcStats._fields_}
        if pixels is None:
            for k in ("rgb_lin", "rgb8"):
                out[k] = out[k].reshape(height, width, 3)
            for k in ("hit_count", "prim0", "t0"):
                out[k] = out[k].reshape(height, width)
        return out


def camera_basis(eye, look_at, up):
    u, v, w = (np.zeros(3, np.float32) for _ in range(3))
    lib().orc_camera_basis(_ptr(_f32(eye)), _ptr(_f32(look_at)), _ptr(_f32(up)), _ptr(u), _ptr(v), _ptr(w))
    return u, v, w


def light_sample(tri, a, b):
    out = np.zeros(3, np.float32)
    lib().orc_light_sample(_ptr(_f32(tri).reshape(9)), float(a), float(b), _ptr(out))
    return out


def box_intersect(mn, mx, o, d):
    return bool(lib().orc_box_intersect(_ptr(_f32(mn)), _ptr(_f32(mx)), _ptr(_f32(o)), _ptr(_f32(d))))


def tri_intersect(v, o, d):
    t = C.c_float(0)
    ok = lib().orc_tri_intersect(_ptr(_f32(v).reshape(9)), _ptr(_f32(o)), _ptr(_f32(d)), C.byref(t))
    return (t.value if ok else None)


def gamma_u8(c):
    return int(lib().orc_gamma_u8(float(np.float32(c))))


# ---- the Sphere primitive (src/tracer/primitives/sphere.rs) ----
def sphere_intersect(center, radius, o, d):
    t = C.c_float(0)
    ok = lib().orc_sphere_intersect(_ptr(_f32(center)), float(np.float32(radius)), _ptr(_f32(o)), _ptr(_f32(d)), C.byref(t))
    return (t.value if ok else None)


def sphere_box(center, radius):
    out = np.zeros(6, np.float32)
    lib().orc_sphere_box(_ptr(_f32(center)), float(np.float32(radius)), _ptr(out))
    return out


def sphere_normal(center, p):
    out = np.zeros(3, np.float32)
    lib().orc_sphere_normal(_ptr(_f32(center)), _ptr(_f32(p)), _ptr(out))
    return out


def spheres_closest(xyzr, origins, dirs):
    """closest sphere per ray -> (idx int32 [-1 = none], t float32, normal float32[n,3])"""
    xyzr = np.ascontiguousarray(xyzr, np.float32).reshape(-1, 4)
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3); d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    idx = np.full(len(o), -1, np.int32); t = np.zeros(len(o), np.float32); nrm = np.zeros((len(o), 3), np.float32)
    L = lib()
    for k in range(len(o)):
        tk = C.c_float(0)
        idx[k] = L.orc_spheres_closest(_ptr(xyzr), len(xyzr), _ptr(o[k]), _ptr(d[k]), C.byref(tk), _ptr(nrm[k]))
        t[k] = tk.value
    return idx, t, nrm
