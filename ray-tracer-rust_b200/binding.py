"""ctypes binding of librtb200.so -- the Python mirror of the reference's entry points.

Names follow the reference: `import_obj` (src/main.rs:114), `create_ground` (src/main.rs:102), `Scene`
(src/tracer/utils/scene.rs:6), `render` (src/main.rs:242), `render_pixel` (src/main.rs:180).  Everything that
computes goes through the C ABI of include/rtb200.h into the CUDA library; there is no Python or CPU
fallback -- if the library is missing or no CUDA device is usable, calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTB200_LIB") or os.path.join(_HERE, "librtb200.so")   # override: kernel-variant experiments

RTB_FLAG_COUNTERS = 1
RTB_FLAG_PROFILE = 2
RTB_FLAG_FRAME_ROWS = 4
RTB_FLAG_GATHER_PEER = 8
COLOR_STATIC_MAX = 1 << 20
RTB_TIE_REFERENCE, RTB_TIE_CALLER, RTB_TIE_INDEX = 0, 1, 2


class RtbError(RuntimeError):
    pass


class SceneDesc(C.Structure):
    _fields_ = [("n_tris", C.c_uint32), ("verts", C.c_void_p), ("colors", C.c_void_p), ("tie_rank", C.c_void_p),
                ("light", C.c_float * 9), ("device", C.c_int32)]


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("nb_ray", C.c_uint32), ("nb_light_sample", C.c_uint32),
                ("eye", C.c_float * 3), ("look_at", C.c_float * 3), ("up", C.c_float * 3), ("distance", C.c_float),
                ("band_rows", C.c_uint32), ("rank", C.c_uint32), ("world", C.c_uint32), ("flags", C.c_uint32)]


class Stats(C.Structure):
    _fields_ = [("primary_rays", C.c_uint64), ("primary_hits", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("shadow_occluded", C.c_uint64), ("box_tests", C.c_uint64), ("tri_tests", C.c_uint64),
                ("kernel_ms", C.c_float), ("launches", C.c_uint32), ("primary_ms", C.c_float), ("shadow_ms", C.c_float),
                ("shade_ms", C.c_float), ("queued_hits", C.c_uint32), ("queued_groups", C.c_uint32), ("queued_walks", C.c_uint32),
                ("grid_cells", C.c_uint32), ("grid_entries", C.c_uint32), ("grid_big", C.c_uint32), ("grid_build_ms", C.c_float),
                ("grid_rebuilds", C.c_uint32), ("grid_redo", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SceneInfo(C.Structure):
    _fields_ = [("n_tris", C.c_uint32), ("n_nodes", C.c_uint32), ("max_depth", C.c_uint32), ("ploc_iterations", C.c_uint32),
                ("sah_cost", C.c_float), ("build_ms", C.c_float), ("tie_mode", C.c_uint32)]


class SceneOptions(C.Structure):
    """rtb_scene_options: the tunables that were environment variables (0 = default)."""
    _fields_ = [("struct_size", C.c_uint32), ("shadow_mode", C.c_int32), ("primary_mode", C.c_int32), ("term_mode", C.c_int32),
                ("chunk_rays", C.c_uint64), ("term_bytes", C.c_uint64), ("grid_entry_cap", C.c_uint64),
                ("grid_cell_factor", C.c_float), ("cam_cell_px", C.c_float), ("cdiv_off", C.c_int32), ("trace_chunks", C.c_int32)]


class ImportStats(C.Structure):
    _fields_ = [("read_s", C.c_double), ("parse_s", C.c_double), ("alloc_s", C.c_double), ("file_bytes", C.c_uint64), ("threads", C.c_uint32),
                ("n_vertices", C.c_uint32)]


class CameraPose(C.Structure):
    _fields_ = [("eye", C.c_float * 3), ("look_at", C.c_float * 3), ("up", C.c_float * 3), ("distance", C.c_float)]


class SequenceStats(C.Structure):
    _fields_ = [("seconds", C.c_double), ("device_ms", C.c_float), ("frames", C.c_uint32), ("cam_grid_redo", C.c_uint32),
                ("primary_rays", C.c_uint64), ("shadow_rays", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


EXPORTS = ["rtb_import_obj_parallel_ex", "rtb_spheres_closest_hit", "rtb_spheres_bounds", "rtb_gen_random_spheres", "rtb_gen_random_triangles", "rtb_import_obj_parallel", "rtb_free_pinned", "rtb_color_to_u8_static", "rtb_scene_create", "rtb_scene_create_ex", "rtb_scene_set_samples_async", "rtb_render_multi", "rtb_scenes_set_samples",
           "rtb_render_sequence", "rtb_render_sequence_multi", "rtb_measure_peaks", "rtb_scene_destroy", "rtb_scene_get_info", "rtb_scene_set_samples", "rtb_scene_set_materials", "rtb_render", "rtb_render_begin", "rtb_render_end",
           "rtb_render_device", "rtb_render_pixels", "rtb_color_to_u8", "rtb_scene_get_tie_ranks", "rtb_scene_get_nodes", "rtb_selftest_division", "rtb_scene_get_counters", "rtb_owned_rows", "rtb_deinterleave_bands", "rtb_peer_alloc", "rtb_peer_open", "rtb_peer_close", "rtb_import_obj", "rtb_import_obj_ex",
           "rtb_free", "rtb_fill_samples", "rtb_write_png", "rtb_last_error", "rtb_version"]

_lib = None


def lib():
    """Loads the CUDA library; fails loudly when it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtbError("librtb200.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "or `make -C ray-tracer-rust_b200`")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.rtb_scene_create.argtypes = [vp, vp]
        L.rtb_scene_create_ex.argtypes = [vp, vp, vp]
        L.rtb_scene_set_samples_async.argtypes = [vp, vp, C.c_uint32]
        L.rtb_scenes_set_samples.argtypes = [vp, C.c_uint32, vp, C.c_uint32]
        L.rtb_render_multi.argtypes = [vp, C.c_uint32, vp, vp, vp]
        L.rtb_render_sequence.argtypes = [vp, vp, vp, C.c_uint32, vp, C.c_char_p, vp, vp, vp]
        L.rtb_render_sequence_multi.argtypes = [vp, C.c_uint32, vp, vp, C.c_uint32, vp, C.c_char_p, vp]
        L.rtb_measure_peaks.argtypes = [C.c_int32, vp, vp, vp]
        L.rtb_scene_destroy.argtypes = [vp]; L.rtb_scene_destroy.restype = None
        L.rtb_scene_get_info.argtypes = [vp, vp]
        L.rtb_scene_set_samples.argtypes = [vp, vp, C.c_uint32]
        L.rtb_scene_set_materials.argtypes = [vp, vp]
        L.rtb_render.argtypes = [vp, vp, vp, vp]
        L.rtb_render_begin.argtypes = [vp, vp, vp]
        L.rtb_render_end.argtypes = [vp, vp]
        L.rtb_render_device.argtypes = [vp, vp, vp, vp]
        L.rtb_render_pixels.argtypes = [vp, vp, vp, C.c_uint64, vp, vp, vp, vp, vp, vp]
        L.rtb_color_to_u8.argtypes = [vp, vp, C.c_uint64, vp]
        L.rtb_color_to_u8_static.argtypes = [vp, vp, C.c_uint64, vp]
        L.rtb_scene_get_tie_ranks.argtypes = [vp, vp]
        L.rtb_scene_get_nodes.argtypes = [vp, vp]
        L.rtb_scene_get_counters.argtypes = [vp, vp]
        L.rtb_selftest_division.argtypes = [vp, C.c_uint64, C.c_uint64, vp]
        L.rtb_owned_rows.argtypes = [C.c_uint32] * 4; L.rtb_owned_rows.restype = C.c_uint32
        L.rtb_deinterleave_bands.argtypes = [C.c_int32, vp, C.c_uint64, vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, vp]
        L.rtb_peer_alloc.argtypes = [C.c_int32, C.c_uint64, C.POINTER(vp), C.c_char_p]
        L.rtb_peer_open.argtypes = [C.c_int32, C.c_char_p, C.POINTER(vp)]
        L.rtb_peer_close.argtypes = [C.c_int32, vp, C.c_int32]
        L.rtb_import_obj.argtypes = [C.c_char_p, vp, vp]
        L.rtb_import_obj_ex.argtypes = [C.c_char_p, C.c_uint32, vp, vp, vp, vp]
        L.rtb_free.argtypes = [vp]; L.rtb_free.restype = None
        L.rtb_free_pinned.argtypes = [vp]; L.rtb_free_pinned.restype = None
        L.rtb_import_obj_parallel.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, vp, vp, vp]
        L.rtb_import_obj_parallel_ex.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, vp, vp, vp, vp, vp]
        L.rtb_spheres_closest_hit.argtypes = [C.c_int32, vp, C.c_uint32, vp, vp, C.c_uint64, vp, vp, vp]
        L.rtb_spheres_bounds.argtypes = [vp, C.c_uint32, vp]; L.rtb_spheres_bounds.restype = None
        L.rtb_gen_random_spheres.argtypes = [C.c_uint64, vp, vp]; L.rtb_gen_random_spheres.restype = C.c_uint32
        L.rtb_gen_random_triangles.argtypes = [C.c_uint64, vp, vp]; L.rtb_gen_random_triangles.restype = C.c_uint32
        L.rtb_fill_samples.argtypes = [C.c_uint64, C.c_uint32, vp]; L.rtb_fill_samples.restype = None
        L.rtb_write_png.argtypes = [C.c_char_p, vp, C.c_uint32, C.c_uint32]
        L.rtb_last_error.restype = C.c_char_p
        L.rtb_version.restype = C.c_char_p
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise RtbError("rtb200 error %d: %s" % (rc, lib().rtb_last_error().decode("utf-8", "replace")))


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


# ---- host-side entry points kept from main.rs -------------------------------------------------

def import_obj(path):
    """import_obj (src/main.rs:114-149) -> (n,9) float32.  Like the reference, an unreadable path prints
    'Not a valid path: ...' and yields no triangles."""
    verts = C.c_void_p()
    n = C.c_uint32(0)
    rc = lib().rtb_import_obj(os.fsencode(path), C.byref(verts), C.byref(n))
    if rc == -4 and n.value == 0 and lib().rtb_last_error().startswith(b"Not a valid path"):
        print(lib().rtb_last_error().decode())
        return np.zeros((0, 9), np.float32)
    _check(rc)
    if n.value == 0:
        return np.zeros((0, 9), np.float32)
    out = np.ctypeslib.as_array(C.cast(verts, C.POINTER(C.c_float)), shape=(n.value, 9)).copy()
    lib().rtb_free(verts)
    return out


RTB_OBJ_MTL = 1


def import_obj_ex(path, mtl=True):
    """Robust OBJ + MTL import (rtb_import_obj_ex; an extension, the reference skips materials, src/main.rs:130,137) ->
    (verts (n,9), kd (n,3), ks_ns (n,4)) float32.  mtl=False: every triangle white as in the reference."""
    verts, kd, ksns = C.c_void_p(), C.c_void_p(), C.c_void_p()
    n = C.c_uint32(0)
    _check(lib().rtb_import_obj_ex(os.fsencode(path), RTB_OBJ_MTL if mtl else 0, C.byref(verts), C.byref(kd), C.byref(ksns), C.byref(n)))
    if n.value == 0:
        return np.zeros((0, 9), np.float32), np.zeros((0, 3), np.float32), np.zeros((0, 4), np.float32)
    out = []
    for ptr, w in ((verts, 9), (kd, 3), (ksns, 4)):
        out.append(np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(n.value, w)).copy())
        lib().rtb_free(ptr)
    return tuple(out)


RTB_OBJ_PINNED = 2


class _HostArrayOwner:
    """Keeps a library-allocated host array alive for the numpy view over it; frees it with the matching call."""

    def __init__(self, ptr, pinned):
        self.ptr, self.pinned = ptr, pinned

    def __del__(self):
        if self.ptr:
            (lib().rtb_free_pinned if self.pinned else lib().rtb_free)(self.ptr)
            self.ptr = None


def _owned_view(ptr, n, width, pinned):
    owner = _HostArrayOwner(ptr, pinned)
    buf = (C.c_float * (n * width)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=np.float32).reshape(n, width).view(_OwnedArray)
    arr._owner = owner
    return arr


def import_obj_parallel(path, n_threads=0, pinned=False, mtl=False):
    """rtb_import_obj_parallel[_ex] -> ((n, 9) float32 view of the library's array (page-locked if pinned), stats dict), or with
    mtl=True (verts, kd (n, 3), ks_ns (n, 4), stats).  The views keep the allocations alive; they are freed with the last reference."""
    verts, kd, ks = C.c_void_p(), C.c_void_p(), C.c_void_p()
    n = C.c_uint32(0)
    st = ImportStats()
    flags = (RTB_OBJ_PINNED if pinned else 0) | (RTB_OBJ_MTL if mtl else 0)
    if mtl:
        _check(lib().rtb_import_obj_parallel_ex(os.fsencode(path), n_threads, flags, C.byref(verts), C.byref(kd), C.byref(ks), C.byref(n), C.byref(st)))
    else:
        _check(lib().rtb_import_obj_parallel(os.fsencode(path), n_threads, flags, C.byref(verts), C.byref(n), C.byref(st)))
    stats = {k: getattr(st, k) for k, _ in ImportStats._fields_}
    if n.value == 0:
        z = np.zeros((0, 9), np.float32)
        return (z, np.zeros((0, 3), np.float32), np.zeros((0, 4), np.float32), stats) if mtl else (z, stats)
    arr = _owned_view(verts.value, n.value, 9, pinned)
    if mtl:
        return arr, _owned_view(kd.value, n.value, 3, pinned), _owned_view(ks.value, n.value, 4, pinned), stats
    return arr, stats


class _OwnedArray(np.ndarray):
    """ndarray that carries a reference to the owner of its memory (views of it keep it through .base)."""
    _owner = None


def create_ground():
    """create_ground (src/main.rs:102-111): one grey triangle."""
    return (np.array([[-10000.0, 0.0, -10000.0, 10000.0, 0.0, -10000.0, 0.0, 0.0, 10000.0]], np.float32),
            np.array([[0.5, 0.5, 0.5]], np.float32))


DEFAULT_LIGHT = np.array([-10.0, 300.0, -10.0, 10.0, 300.0, -10.0, 0.0, 300.0, 0.0], np.float32)   # src/main.rs:337-347
DEFAULT_CAMERA = dict(eye=(0.0, 100.0, 200.0), look_at=(0.0, 0.0, -100000.0), up=(0.0, 1.0, 0.0), distance=288.0)  # :353-356


def fill_samples(n_pairs=2000000, seed=0x5EED0001):
    out = np.empty((n_pairs, 2), np.float32)
    lib().rtb_fill_samples(C.c_uint64(seed), n_pairs, _ptr(out))
    return out


def write_png(path, rgb8):
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    h, w, _ = rgb8.shape
    _check(lib().rtb_write_png(os.fsencode(path), _ptr(rgb8), w, h))


def owned_rows(height, band_rows, rank, world):
    return int(lib().rtb_owned_rows(height, band_rows, rank, world))


def owned_row_indices(height, band_rows, rank, world):
    """Global row numbers, in storage order, of the rows (rank, world) owns."""
    if world <= 1:
        return np.arange(height)
    rows = [np.arange(b * band_rows, min(height, (b + 1) * band_rows))
            for b in range(rank, (height + band_rows - 1) // band_rows, world)]
    return np.concatenate(rows) if rows else np.zeros(0, np.int64)


# ---- the scene handle ---------------------------------------------------------------------------

class Scene:
    """Device-resident Scene{light, camera, bvh} (src/tracer/utils/scene.rs:6-12).  Primitives are given in the
    reference's order: OBJ triangles in argv order, ground last (src/main.rs:324-335)."""

    def __init__(self, verts, colors, light=DEFAULT_LIGHT, device=0, tie_rank=None, **options):
        self.verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        self.colors = None if colors is None else np.ascontiguousarray(colors, np.float32).reshape(-1, 3)   # None: all white (src/main.rs:143)
        if self.colors is not None and self.verts.shape[0] != self.colors.shape[0]:
            raise ValueError("verts/colors length mismatch")
        d = SceneDesc()
        d.n_tris = self.verts.shape[0]
        d.verts = self.verts.ctypes.data if d.n_tris else None
        d.colors = self.colors.ctypes.data if (d.n_tris and self.colors is not None) else None
        self._rank = None if tie_rank is None else np.ascontiguousarray(tie_rank, np.int32)
        d.tie_rank = self._rank.ctypes.data if self._rank is not None else None
        d.light[:] = [float(x) for x in np.asarray(light, np.float32).reshape(9)]
        d.device = device
        self.device = device
        h = C.c_void_p()
        if options:                                  # rtb_scene_options fields by name (shadow_mode=1, chunk_rays=6000, ...)
            o = SceneOptions()
            o.struct_size = C.sizeof(SceneOptions)
            for k, v in options.items():
                if k not in dict(SceneOptions._fields_):
                    raise TypeError("unknown scene option %r" % k)
                setattr(o, k, v)
            _check(lib().rtb_scene_create_ex(C.byref(d), C.byref(o), C.byref(h)))
        else:
            _check(lib().rtb_scene_create(C.byref(d), C.byref(h)))
        self._h = h
        self._samples = None

    def close(self):
        if getattr(self, "_h", None):
            lib().rtb_scene_destroy(self._h)
            self._h = None

    __del__ = close

    def info(self):
        i = SceneInfo()
        _check(lib().rtb_scene_get_info(self._h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in SceneInfo._fields_}

    def set_samples(self, samples):
        s = np.ascontiguousarray(samples, np.float32).reshape(-1, 2)
        _check(lib().rtb_scene_set_samples(self._h, _ptr(s), s.shape[0]))
        self._samples = s

    def set_samples_async(self, samples):
        """rtb_scene_set_samples_async: `samples` must be pinned host memory that stays valid until the next render returns."""
        s = np.ascontiguousarray(samples, np.float32).reshape(-1, 2)
        _check(lib().rtb_scene_set_samples_async(self._h, _ptr(s), s.shape[0]))
        self._samples_prev = getattr(self, "_samples", None)      # a frame in flight may still be reading the previous table
        self._samples = s

    def render_sequence(self, poses, frames=None, png_pattern=None, **kw):
        """rtb_render_sequence: poses = [(eye, look_at, up[, distance])]; returns (frames (n, rows, width, 3) uint8, stats)."""
        p = self.params(**kw)
        n = len(poses)
        arr = (CameraPose * max(n, 1))()
        for k, c in enumerate(poses):
            arr[k].eye[:] = [float(x) for x in c[0]]; arr[k].look_at[:] = [float(x) for x in c[1]]; arr[k].up[:] = [float(x) for x in c[2]]
            arr[k].distance = float(c[3]) if len(c) > 3 else p.distance
        rows = owned_rows(p.height, p.band_rows, p.rank, max(1, p.world))
        if frames is None and png_pattern is None:
            frames = np.empty((n, rows, p.width, 3), np.uint8)
        st = SequenceStats()
        _check(lib().rtb_render_sequence(self._h, C.byref(p), arr, n, _ptr(frames), os.fsencode(png_pattern) if png_pattern else None,
                                         None, None, C.byref(st)))
        return frames, st.as_dict()

    def set_materials(self, ks_ns):
        """Phong extension (rtb_scene_set_materials): (n_tris,4) {Ks.rgb, Ns} or None for the reference's Lambert."""
        if ks_ns is None:
            _check(lib().rtb_scene_set_materials(self._h, None))
            return
        k = np.ascontiguousarray(ks_ns, np.float32).reshape(-1, 4)
        if k.shape[0] != self.info()["n_tris"]:
            raise ValueError("ks_ns needs one row per primitive")
        _check(lib().rtb_scene_set_materials(self._h, _ptr(k)))

    @staticmethod
    def params(width=1920, height=1080, nb_ray=1, nb_light_sample=100, eye=DEFAULT_CAMERA["eye"],
               look_at=DEFAULT_CAMERA["look_at"], up=DEFAULT_CAMERA["up"], distance=DEFAULT_CAMERA["distance"],
               band_rows=8, rank=0, world=1, counters=False, profile=False, frame_rows=False):
        p = RenderParams()
        p.width, p.height, p.nb_ray, p.nb_light_sample = width, height, nb_ray, nb_light_sample
        p.eye[:] = [float(x) for x in eye]; p.look_at[:] = [float(x) for x in look_at]; p.up[:] = [float(x) for x in up]
        p.distance = float(distance)
        p.band_rows, p.rank, p.world = band_rows, rank, world
        p.flags = (RTB_FLAG_COUNTERS if counters else 0) | (RTB_FLAG_PROFILE if profile else 0) | (RTB_FLAG_FRAME_ROWS if frame_rows else 0)
        return p

    def render(self, out=None, **kw):
        """render(scene) (src/main.rs:242-310) minus the PNG save -> ((owned_rows, width, 3) uint8, stats)."""
        p = self.params(**kw)
        rows = owned_rows(p.height, p.band_rows, p.rank, max(1, p.world))
        if out is None:
            out = np.empty((rows, p.width, 3), np.uint8)
        st = Stats()
        _check(lib().rtb_render(self._h, C.byref(p), _ptr(out), C.byref(st)))
        return out, st.as_dict()

    def render_begin(self, out, **kw):
        """First half of render(): enqueues the frame into the PINNED array `out` and returns (rtb_render_begin).  A
        set_samples_async() before render_end() hands over the next frame's table."""
        p = self.params(**kw)
        self._inflight = (p, out)                       # both must outlive the frame
        _check(lib().rtb_render_begin(self._h, C.byref(p), _ptr(out)))

    def render_end(self):
        """Second half: waits for the frame begun with render_begin -> (out, stats)."""
        st = Stats()
        _check(lib().rtb_render_end(self._h, C.byref(st)))
        p, out = self._inflight
        self._inflight = None
        return out, st.as_dict()

    def color_to_u8(self, linear):
        """Color::to_rgba channels (color.rs:26-33) on the device."""
        lin = np.ascontiguousarray(linear, np.float32)
        out = np.zeros(lin.shape, np.uint8)
        _check(lib().rtb_color_to_u8(self._h, _ptr(lin), lin.size, _ptr(out)))
        return out

    def color_to_u8_static(self, linear):
        """rtb_color_to_u8_static: the same conversion with no allocation (static device scratch, at most COLOR_STATIC_MAX values)."""
        lin = np.ascontiguousarray(linear, np.float32)
        out = np.zeros(lin.shape, np.uint8)
        _check(lib().rtb_color_to_u8_static(self._h, _ptr(lin), lin.size, _ptr(out)))
        return out

    def nodes(self):
        """(n_nodes, 16) float32 view of the BVH + (n_nodes, 2) int32 children (see rtb_scene_get_nodes)."""
        n = self.info()["n_nodes"]
        out = np.zeros((n, 16), np.float32)
        _check(lib().rtb_scene_get_nodes(self._h, _ptr(out)))
        return out, out[:, 12:14].copy().view(np.int32)

    def counters(self):
        out = np.zeros(64, np.uint64)                # the library writes 16 values (diagnostic builds: up to 64)
        _check(lib().rtb_scene_get_counters(self._h, _ptr(out)))
        return out

    def selftest_division(self, n=1 << 26, seed=1):
        """Mismatches between the kernels' shared-reciprocal division and IEEE `/` over n random operand sets (must be 0)."""
        bad = C.c_uint64(0)
        _check(lib().rtb_selftest_division(self._h, seed, n, C.byref(bad)))
        return bad.value

    def tie_ranks(self):
        out = np.zeros(self.verts.shape[0], np.int32)
        _check(lib().rtb_scene_get_tie_ranks(self._h, _ptr(out)))
        return out

    def render_device(self, d_ptr, stream=0, **kw):
        """Asynchronous render into device memory (d_ptr: int device address, stream: cudaStream_t as int)."""
        p = self.params(**kw)
        _check(lib().rtb_render_device(self._h, C.byref(p), C.c_void_p(d_ptr), C.c_void_p(stream)))

    def render_pixels(self, pixels, **kw):
        """render_pixel (src/main.rs:180-240) for a list of (px,py) -> dict of per-pixel arrays + stats."""
        p = self.params(**kw)
        pixels = np.ascontiguousarray(pixels, np.uint32).reshape(-1, 2)
        m = pixels.shape[0]
        out = dict(rgb_lin=np.zeros((m, 3), np.float32), rgb8=np.zeros((m, 3), np.uint8), hit_count=np.zeros(m, np.uint8),
                   prim0=np.zeros(m, np.int32), t0=np.zeros(m, np.float32))
        st = Stats()
        _check(lib().rtb_render_pixels(self._h, C.byref(p), _ptr(pixels), m, _ptr(out["rgb_lin"]), _ptr(out["rgb8"]),
                                       _ptr(out["hit_count"]), _ptr(out["prim0"]), _ptr(out["t0"]), C.byref(st)))
        out["stats"] = st.as_dict()
        return out


def _scene_array(scenes):
    arr = (C.c_void_p * len(scenes))()
    for i, sc in enumerate(scenes):
        arr[i] = sc._h
    return arr


def set_samples_multi(scenes, samples):
    """rtb_scenes_set_samples: the same table on every scene, uploaded by one host thread per device (pinned memory: async)."""
    s = np.ascontiguousarray(samples, np.float32).reshape(-1, 2)
    _check(lib().rtb_scenes_set_samples(_scene_array(scenes), len(scenes), _ptr(s), s.shape[0]))
    for sc in scenes:
        sc._samples = s


def render_multi(scenes, out=None, gather_peer=False, **kw):
    """rtb_render_multi: one frame over len(scenes) GPUs (scenes[i] lives on device i) -> ((height, width, 3) uint8, stats)."""
    p = Scene.params(**kw)
    if gather_peer:
        p.flags |= RTB_FLAG_GATHER_PEER
    if out is None:
        out = np.empty((p.height, p.width, 3), np.uint8)
    st = Stats()
    _check(lib().rtb_render_multi(_scene_array(scenes), len(scenes), C.byref(p), _ptr(out), C.byref(st)))
    return out, st.as_dict()


def render_sequence_multi(scenes, poses, frames=None, png_pattern=None, **kw):
    """rtb_render_sequence_multi: pose k on scenes[k % n] (frame-parallel) -> ((n, height, width, 3) uint8, stats)."""
    p = Scene.params(**kw)
    n = len(poses)
    arr = (CameraPose * max(n, 1))()
    for k, c in enumerate(poses):
        arr[k].eye[:] = [float(x) for x in c[0]]; arr[k].look_at[:] = [float(x) for x in c[1]]; arr[k].up[:] = [float(x) for x in c[2]]
        arr[k].distance = float(c[3]) if len(c) > 3 else p.distance
    if frames is None and png_pattern is None:
        frames = np.empty((n, p.height, p.width, 3), np.uint8)
    st = SequenceStats()
    _check(lib().rtb_render_sequence_multi(_scene_array(scenes), len(scenes), C.byref(p), arr, n, _ptr(frames),
                                           os.fsencode(png_pattern) if png_pattern else None, C.byref(st)))
    return frames, st.as_dict()


# ---- the Sphere primitive and the random scene generators (dead code in the reference; see include/rtb200.h) ----
def spheres_closest_hit(spheres_xyzr, origins, dirs, device=0):
    """rtb_spheres_closest_hit -> (idx int32 [-1 = none], t float32, normal float32 [n, 3])"""
    sp = np.ascontiguousarray(spheres_xyzr, np.float32).reshape(-1, 4)
    o = np.ascontiguousarray(origins, np.float32).reshape(-1, 3); d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    t = np.zeros(len(o), np.float32); idx = np.zeros(len(o), np.int32); nrm = np.zeros((len(o), 3), np.float32)
    _check(lib().rtb_spheres_closest_hit(device, _ptr(sp), len(sp), _ptr(o), _ptr(d), len(o), _ptr(t), _ptr(idx), _ptr(nrm)))
    return idx, t, nrm


def spheres_bounds(spheres_xyzr):
    sp = np.ascontiguousarray(spheres_xyzr, np.float32).reshape(-1, 4)
    out = np.zeros((len(sp), 6), np.float32)
    lib().rtb_spheres_bounds(_ptr(sp), len(sp), _ptr(out))
    return out


def gen_random_spheres(seed):
    """gen_random_spheres (src/main.rs:43-67) from an explicit seed -> (spheres [1, 4] {x, y, z, radius}, colors [1, 3])"""
    sp, col = np.zeros((1, 4), np.float32), np.zeros((1, 3), np.float32)
    n = lib().rtb_gen_random_spheres(int(seed), _ptr(sp), _ptr(col))
    return sp[:n], col[:n]


def gen_random_triangles(seed):
    """gen_random_triangles (src/main.rs:69-99) from an explicit seed -> (verts [10, 9], colors [10, 3])"""
    v, col = np.zeros((10, 9), np.float32), np.zeros((10, 3), np.float32)
    n = lib().rtb_gen_random_triangles(int(seed), _ptr(v), _ptr(col))
    return v[:n], col[:n]


def measure_peaks(device=0, hbm=True):
    """rtb_measure_peaks -> dict(fp32_tflops, l2_read_gbs, hbm_read_gbs): the roofline denominators, measured."""
    f, l, h = C.c_float(0), C.c_float(0), C.c_float(0)
    _check(lib().rtb_measure_peaks(device, C.byref(f), C.byref(l), C.byref(h) if hbm else None))
    return {"fp32_tflops": f.value, "l2_read_gbs": l.value, "hbm_read_gbs": h.value if hbm else None}


def deinterleave_bands(device, d_staged, stride_bytes, d_frame, width, height, band_rows, world, stream=0):
    _check(lib().rtb_deinterleave_bands(device, C.c_void_p(d_staged), stride_bytes, C.c_void_p(d_frame), width, height,
                                        band_rows, world, C.c_void_p(stream)))


def peer_alloc(device, nbytes):
    """Rank 0: device buffer + its 64-byte CUDA IPC handle (rtb_peer_alloc)."""
    ptr = C.c_void_p()
    handle = C.create_string_buffer(64)
    _check(lib().rtb_peer_alloc(device, nbytes, C.byref(ptr), handle))
    return ptr.value, handle.raw


def peer_open(device, handle):
    """Other ranks: map rank 0's buffer into this process (peer access over NVLink)."""
    ptr = C.c_void_p()
    _check(lib().rtb_peer_open(device, bytes(handle), C.byref(ptr)))
    return ptr.value


def peer_close(device, ptr, owner):
    _check(lib().rtb_peer_close(device, C.c_void_p(ptr), 1 if owner else 0))
