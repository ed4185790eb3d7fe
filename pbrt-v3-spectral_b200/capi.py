"""ctypes binding of include/sptx.h (the C ABI of libsptx.so).

This is harness glue for tests and benchmarks: it loads the in-tree
``libsptx.so`` and exposes the entry points one-to-one.  There is no Python or
CPU implementation behind it -- if the library is missing, or no CUDA device is
present, calls fail loudly (north star: "no CPU fallback").
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# SPTX_LIB_PATH: a diagnosis build of the same library (e.g. -DSPTX_STAGE_CLOCKS), tools/ only
LIB_PATH = os.environ.get("SPTX_LIB_PATH") or os.path.join(HERE, "libsptx.so")

NBINS = 31
STRIDE = 32

# Every symbol include/sptx.h declares (tests check the library exports them all).
SYMBOLS = [
    "sptx_abi_version", "sptx_create", "sptx_destroy", "sptx_last_error",
    "sptx_upload_scene", "sptx_intersect_batch", "sptx_intersect_batch_device",
    "sptx_intersect_batch_noalloc", "sptx_measure_l2", "sptx_bvh_build_ms",
    "sptx_render", "sptx_render_tiles", "sptx_film_clear", "sptx_film_device_ptr",
    "sptx_film_export", "sptx_film_merge_peers", "sptx_film_reduce_scatter", "sptx_film_merge_contexts",
    "sptx_film_reduce",
    "sptx_film_download", "sptx_film_pack_dat", "sptx_write_dat",
    "sptx_get_counters", "sptx_reset_counters", "sptx_set_wavefront_size", "sptx_set_option", "sptx_last_batch_sorted", "sptx_light_table_rows",
    "sptx_sampler_eval", "sptx_camera_rays", "sptx_debug_radiance",
    "sptx_scene_save", "sptx_scene_load", "sptx_scene_free", "sptx_io_last_error",
]

IPC_HANDLE_BYTES = 64
NOALLOC_MAX_RAYS = 65536   # SPTX_NOALLOC_MAX_RAYS
RAY_DTYPE = np.dtype([("o", "<f4", 3), ("t_max", "<f4"), ("d", "<f4", 3), ("time", "<f4")])
HIT_DTYPE = np.dtype([("prim", "<i4"), ("instance", "<i4"), ("t", "<f4"), ("b0", "<f4"),
                      ("b1", "<f4"), ("b2", "<f4"), ("n_nodes", "<i4"), ("n_tests", "<i4")])
assert RAY_DTYPE.itemsize == 32 and HIT_DTYPE.itemsize == 32


class Counters(C.Structure):
    _fields_ = [("camera_rays", C.c_uint64), ("closest_hit_rays", C.c_uint64),
                ("shadow_rays", C.c_uint64), ("path_vertices", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("render_ms", C.c_double),
                ("traverse_ms", C.c_double), ("shade_ms", C.c_double), ("film_ms", C.c_double),
                ("kernel_ms", C.c_double * 8), ("kernel_count", C.c_uint64 * 8),
                ("nodes_visited", C.c_uint64 * 2), ("prims_tested", C.c_uint64 * 2)]

KERNEL_NAMES = ["generate", "intersect", "shade", "resolve", "film", "film_spill", "ray_batch"]


class SptxError(RuntimeError):
    pass


_lib = None


def lib():
    """Loads libsptx.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = LIB_PATH
    if os.environ.get("SPTX_LIB"):   # experiments: another build of the same library (Makefile LIB=)
        path = os.path.join(os.path.dirname(LIB_PATH), os.environ["SPTX_LIB"])
    if not os.path.exists(path):
        raise SptxError(
            f"{path} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no fallback implementation)")
    L = C.CDLL(path)
    vp, i32, i64, f32p = C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_float)
    i32p = C.POINTER(C.c_int32)
    L.sptx_abi_version.restype = i32
    L.sptx_create.argtypes = [i32, C.POINTER(vp)]
    L.sptx_destroy.argtypes = [vp]
    L.sptx_destroy.restype = None
    L.sptx_last_error.argtypes = [vp]
    L.sptx_last_error.restype = C.c_char_p
    L.sptx_io_last_error.restype = C.c_char_p
    L.sptx_upload_scene.argtypes = [vp, vp]
    L.sptx_intersect_batch.argtypes = [vp, vp, i64, vp, i32, i32]
    L.sptx_intersect_batch_device.argtypes = [vp, vp, i64, vp, i32, i32, f32p]
    L.sptx_intersect_batch_noalloc.argtypes = [vp, vp, i64, vp, i32]
    L.sptx_measure_l2.argtypes = [vp, i64, i32, C.POINTER(C.c_double)]
    L.sptx_bvh_build_ms.argtypes = [vp]
    L.sptx_bvh_build_ms.restype = C.c_double
    L.sptx_render.argtypes = [vp, i32p, i32, i32]
    L.sptx_render_tiles.argtypes = [vp, i32, i32, i32, i32]
    L.sptx_film_clear.argtypes = [vp]
    L.sptx_film_device_ptr.argtypes = [vp, C.POINTER(vp), C.POINTER(i64)]
    L.sptx_film_download.argtypes = [vp, vp, vp]
    L.sptx_film_pack_dat.argtypes = [vp, vp]
    L.sptx_film_export.argtypes = [vp, vp]
    L.sptx_film_merge_peers.argtypes = [vp, vp, i32, i32, vp]
    L.sptx_film_reduce_scatter.argtypes = [vp, vp, i32, i32, i32]
    L.sptx_film_merge_contexts.argtypes = [C.POINTER(vp), i32, i32]
    L.sptx_film_reduce.argtypes = [vp, vp, i32]
    L.sptx_write_dat.argtypes = [C.c_char_p, i32, i32, vp]
    L.sptx_get_counters.argtypes = [vp, C.POINTER(Counters)]
    L.sptx_reset_counters.argtypes = [vp]
    L.sptx_set_wavefront_size.argtypes = [vp, i64]
    L.sptx_set_option.argtypes = [vp, C.c_char_p, i64]
    L.sptx_last_batch_sorted.argtypes = [vp]
    L.sptx_light_table_rows.argtypes = [vp]
    L.sptx_light_table_rows.restype = i64
    L.sptx_sampler_eval.argtypes = [vp, vp, vp, vp, vp, i64, vp]
    L.sptx_camera_rays.argtypes = [vp, vp, i64, C.c_float, vp, vp]
    L.sptx_debug_radiance.argtypes = [vp, vp, vp, vp, i64, vp, vp]
    L.sptx_scene_save.argtypes = [C.c_char_p, vp]
    L.sptx_scene_load.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.sptx_scene_free.argtypes = [vp]
    L.sptx_scene_free.restype = None
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def unpack_blob(path):
    """`path` (a .sptx scene blob); when only its committed .xz form exists (the large configuration
    blobs are kept compressed in git), it is unpacked next to it first.  Returns `path`, or None."""
    if os.path.exists(path):
        return path
    if os.path.exists(path + ".xz"):
        import lzma
        tmp = path + f".tmp{os.getpid()}"
        with lzma.open(path + ".xz", "rb") as src, open(tmp, "wb") as dst:
            while True:
                chunk = src.read(1 << 24)
                if not chunk:
                    break
                dst.write(chunk)
        os.replace(tmp, path)
        return path
    return None


class SceneFile:
    """A flattened scene (host arrays) loaded from a file host/flatten.cpp wrote."""

    def __init__(self, path):
        self.path = path
        self.handle = C.c_void_p()
        rc = lib().sptx_scene_load(path.encode(), C.byref(self.handle))
        if rc != 0:
            raise SptxError(f"sptx_scene_load({path}) -> {rc}: {lib().sptx_io_last_error().decode()}")

    def close(self):
        if self.handle:
            lib().sptx_scene_free(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Context:
    """One GPU context (sptx_ctx).  Mirrors the call sequence of GpuPathIntegrator::Render."""

    def __init__(self, device=0, exact=None):
        """exact: K3's arithmetic -- True = the reference's roundings, False = the fast build of
        K3 (FMA contraction, approximate division / square root), None = the library's default
        (fast unless SPTX_EXACT=1 is in the environment)."""
        self.h = C.c_void_p()
        rc = lib().sptx_create(int(device), C.byref(self.h))
        if rc != 0:
            raise SptxError(f"sptx_create -> {rc}: {lib().sptx_last_error(None).decode()}")
        self.film_shape = None
        env = os.environ.get("SPTX_EXACT", "")
        self.exact = bool(env and env[0] != "0") if exact is None else bool(exact)
        if exact is not None:
            self.set_option("shade_exact", int(self.exact))

    @property
    def shade_arithmetic(self):
        return "exact" if self.exact else "fast"

    def _ck(self, rc, what):
        if rc != 0:
            raise SptxError(f"{what} -> {rc}: {lib().sptx_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            lib().sptx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, scene):
        self._ck(lib().sptx_upload_scene(self.h, scene.handle), "sptx_upload_scene")
        ptr, n = C.c_void_p(), C.c_int64()
        self._ck(lib().sptx_film_device_ptr(self.h, C.byref(ptr), C.byref(n)), "sptx_film_device_ptr")
        self.film_ptr, self.film_floats = ptr.value, n.value

    def set_wavefront_size(self, n):
        self._ck(lib().sptx_set_wavefront_size(self.h, int(n)), "sptx_set_wavefront_size")

    def light_table_rows(self):
        """Voxel rows of the spatial light distribution built by the last upload."""
        return int(lib().sptx_light_table_rows(self.h))

    def last_batch_sorted(self):
        """True when the last ray batch was walked in binned order (set_option("ray_sort", ..))."""
        return bool(lib().sptx_last_batch_sorted(self.h))

    def set_option(self, name, value):
        self._ck(lib().sptx_set_option(self.h, name.encode(), int(value)), "sptx_set_option")

    def intersect(self, rays, any_hit=False, counters=False):
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        self._ck(lib().sptx_intersect_batch(self.h, _ptr(rays), rays.shape[0], _ptr(hits),
                                            int(any_hit), int(counters)), "sptx_intersect_batch")
        return hits

    def intersect_noalloc(self, rays, any_hit=False):
        """sptx_intersect_batch_noalloc: at most NOALLOC_MAX_RAYS rays, fixed device scratch."""
        rays = np.ascontiguousarray(rays, dtype=RAY_DTYPE)
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        self._ck(lib().sptx_intersect_batch_noalloc(self.h, _ptr(rays), rays.shape[0], _ptr(hits),
                                                    int(any_hit)), "sptx_intersect_batch_noalloc")
        return hits

    def bvh_build_ms(self):
        """Device time of the last on-device BVH build (set_option("device_bvh", 1)), else 0."""
        return lib().sptx_bvh_build_ms(self.h)

    def measure_l2(self, nbytes=32 << 20, reps=200):
        """GB/s of 128-bit reads from an L2-resident buffer (sptx_measure_l2)."""
        gbs = C.c_double()
        self._ck(lib().sptx_measure_l2(self.h, int(nbytes), int(reps), C.byref(gbs)), "sptx_measure_l2")
        return gbs.value

    def intersect_device(self, d_rays, n, d_hits, any_hit=False, counters=False):
        ms = C.c_float()
        self._ck(lib().sptx_intersect_batch_device(self.h, d_rays, int(n), d_hits, int(any_hit),
                                                   int(counters), C.byref(ms)),
                 "sptx_intersect_batch_device")
        return ms.value

    def render(self, sample_begin, sample_end, pixel_bounds=None):
        pb = None
        if pixel_bounds is not None:
            pb = (C.c_int32 * 4)(*[int(v) for v in pixel_bounds])
        self._ck(lib().sptx_render(self.h, pb, int(sample_begin), int(sample_end)), "sptx_render")

    def render_tiles(self, part, n_parts, sample_begin, sample_end):
        self._ck(lib().sptx_render_tiles(self.h, int(part), int(n_parts), int(sample_begin),
                                         int(sample_end)), "sptx_render_tiles")

    def film_clear(self):
        self._ck(lib().sptx_film_clear(self.h), "sptx_film_clear")

    def film_download(self, width, height, out=None):
        """contribSum [H][W][31] and filterWeightSum [H][W]; `out` = (contrib, weight) reuses
        caller buffers (float32, C-contiguous) instead of allocating."""
        if out is not None:
            contrib, weight = out
            assert contrib.dtype == np.float32 and contrib.flags.c_contiguous and contrib.size == height * width * NBINS
            assert weight.dtype == np.float32 and weight.flags.c_contiguous and weight.size == height * width
        else:
            contrib = np.empty((height, width, NBINS), dtype=np.float32)
            weight = np.empty((height, width), dtype=np.float32)
        self._ck(lib().sptx_film_download(self.h, _ptr(contrib), _ptr(weight)), "sptx_film_download")
        return contrib, weight

    def film_export(self):
        """64-byte CUDA IPC handle of the device film (for the merging rank)."""
        buf = C.create_string_buffer(IPC_HANDLE_BYTES)
        self._ck(lib().sptx_film_export(self.h, buf), "sptx_film_export")
        return buf.raw

    def film_reduce_scatter(self, handles, rank, dst=0):
        """This rank's slice of the multi-GPU film sum, written into rank dst's film
        (sptx_film_reduce_scatter); every rank calls it between two barriers."""
        buf = b"".join(bytes(h) for h in handles)
        self._ck(lib().sptx_film_reduce_scatter(self.h, buf, len(handles), int(rank), int(dst)),
                 "sptx_film_reduce_scatter")

    def film_reduce_nccl(self, comm, root=0):
        """ncclReduce of the device film onto `root` (sptx_film_reduce); comm = ncclComm_t value."""
        self._ck(lib().sptx_film_reduce(self.h, C.c_void_p(comm), int(root)), "sptx_film_reduce")

    def film_merge_peers(self, handles, self_rank, pack=None):
        """Sums the films of all ranks (rank order) into this rank's film with one kernel reading
        the peers over NVLink; `handles` = list of film_export() results indexed by rank.  With
        pack=(width, height) also returns the .dat planes [31][H][W] float64 of the merged film."""
        blob = b"".join(bytes(h) for h in handles)
        assert len(blob) == IPC_HANDLE_BYTES * len(handles)
        planes = None
        if pack is not None:
            planes = np.empty((NBINS, pack[1], pack[0]), dtype=np.float64)
        self._ck(lib().sptx_film_merge_peers(self.h, blob, len(handles), int(self_rank),
                                             _ptr(planes) if planes is not None else None),
                 "sptx_film_merge_peers")
        return planes

    def film_pack_dat(self, width, height):
        planes = np.zeros((NBINS, height, width), dtype=np.float64)
        self._ck(lib().sptx_film_pack_dat(self.h, _ptr(planes)), "sptx_film_pack_dat")
        return planes

    def counters(self):
        c = Counters()
        self._ck(lib().sptx_get_counters(self.h, C.byref(c)), "sptx_get_counters")
        out = {}
        for k, _ in Counters._fields_:
            v = getattr(c, k)
            out[k] = list(v) if hasattr(v, "__len__") else v
        return out

    def reset_counters(self):
        self._ck(lib().sptx_reset_counters(self.h), "sptx_reset_counters")

    def sampler_eval(self, px, py, sample, dim):
        px, py, sample, dim = [np.ascontiguousarray(a, dtype=np.int32) for a in (px, py, sample, dim)]
        out = np.zeros(px.shape[0], dtype=np.float32)
        self._ck(lib().sptx_sampler_eval(self.h, _ptr(px), _ptr(py), _ptr(sample), _ptr(dim),
                                         px.shape[0], _ptr(out)), "sptx_sampler_eval")
        return out

    def camera_rays(self, cs, wavelength=550.0):
        cs = np.ascontiguousarray(cs, dtype=np.float32).reshape(-1, 5)
        rays = np.zeros(cs.shape[0], dtype=RAY_DTYPE)
        w = np.zeros(cs.shape[0], dtype=np.float32)
        self._ck(lib().sptx_camera_rays(self.h, _ptr(cs), cs.shape[0], float(wavelength), _ptr(rays),
                                        _ptr(w)), "sptx_camera_rays")
        return rays, w

    def debug_radiance(self, px, py, sample):
        px, py, sample = [np.ascontiguousarray(a, dtype=np.int32) for a in (px, py, sample)]
        L = np.zeros((px.shape[0], NBINS), dtype=np.float32)
        pw = np.zeros((px.shape[0], 3), dtype=np.float32)
        self._ck(lib().sptx_debug_radiance(self.h, _ptr(px), _ptr(py), _ptr(sample), px.shape[0],
                                           _ptr(L), _ptr(pw)), "sptx_debug_radiance")
        return L, pw


def merge_contexts(contexts, dst=0):
    """Films of several contexts of THIS process summed into contexts[dst]'s (sptx_film_merge_contexts)."""
    arr = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    rc = lib().sptx_film_merge_contexts(arr, len(contexts), int(dst))
    if rc != 0:
        raise SptxError(f"sptx_film_merge_contexts -> {rc}: {lib().sptx_last_error(contexts[dst].h).decode()}")


def write_dat(path, planes):
    planes = np.ascontiguousarray(planes, dtype=np.float64)
    rc = lib().sptx_write_dat(path.encode(), planes.shape[2], planes.shape[1], _ptr(planes))
    if rc != 0:
        raise SptxError(f"sptx_write_dat -> {rc}")


def read_dat(path):
    """Reads a spectral .dat as Film::WriteImage writes it (reference src/core/film.cpp:267-306)."""
    with open(path, "rb") as f:
        header = f.readline().split()
        w, h, nb = int(header[0]), int(header[1]), int(header[2])
        f.readline()  # "v3 \n"
        data = np.frombuffer(f.read(), dtype="<f8")
    return data.reshape(nb, h, w)
