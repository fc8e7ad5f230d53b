"""gaussian-ray-tracing_b200 -- B200 (sm_100a) software Gaussian ray tracer.

The product is ``csrc/libgrt.so`` (hand-written CUDA behind the C ABI of ``include/grt.h``) plus the C++
host mirror ``host/libgrt_host.so``.  This Python package is only a thin ctypes harness over that C ABI
for tests and ``bench.py``; it contains no arithmetic of the hot path and has no CPU fallback: importing
works without a GPU (so that symbol checks can run), but creating a ``Context`` without one raises.

Import with ``importlib.import_module("gaussian-ray-tracing_b200")`` (the directory name is the repo's
mandated package name and is not a Python identifier).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIBGRT_PATH = os.path.join(_HERE, "csrc", "libgrt.so")
LIBHOST_PATH = os.path.join(_HERE, "host", "libgrt_host.so")

PARTICLE_BYTES = 236  # reference GaussianParticle, src/GaussianData.h:12-20
PARTICLE_FLOATS = 59

MIRROR, NORMAL, GLASS = 0, 1, 2  # MeshType, src/Parameters.h:80-85


class GrtError(RuntimeError):
    """Raised where the reference would throw std::runtime_error (src/Exception.h:31-80)."""


class GrtParams(C.Structure):
    """Byte-identical to the reference ``struct Params`` (src/Parameters.h:42-74), 144 bytes."""

    _fields_ = [
        ("output_buffer", C.c_void_p),
        ("width", C.c_uint32),
        ("height", C.c_uint32),
        ("sh_degree_max", C.c_uint32),
        ("eye", C.c_float * 3),
        ("U", C.c_float * 3),
        ("V", C.c_float * 3),
        ("W", C.c_float * 3),
        ("t_min", C.c_float),
        ("t_max", C.c_float),
        ("minTransmittance", C.c_float),
        ("alpha_min", C.c_float),
        ("handle", C.c_uint64),
        ("d_particles", C.c_void_p),
        ("mesh_handle", C.c_uint64),
        ("mode_fisheye", C.c_uint8),
        ("d_meshes", C.c_void_p),
        ("type", C.c_int32),
        ("traceState", C.c_void_p),
    ]


assert C.sizeof(GrtParams) == 144


class GrtMeshDesc(C.Structure):
    _fields_ = [
        ("vertices", C.POINTER(C.c_float)),
        ("normals", C.POINTER(C.c_float)),
        ("indices", C.POINTER(C.c_uint32)),
        ("vertex_count", C.c_uint32),
        ("index_count", C.c_uint32),
        ("transform", C.c_float * 16),
    ]


class GrtBuildInfo(C.Structure):
    _fields_ = [
        ("n_particles", C.c_uint64),
        ("n_live", C.c_uint64),
        ("n_wide_nodes", C.c_uint32),
        ("wide_depth_levels", C.c_uint32),
        ("n_mesh_triangles", C.c_uint32),
        ("n_mesh_wide_nodes", C.c_uint32),
        ("build_ms", C.c_float),
        ("mesh_build_ms", C.c_float),
        ("scene_lo", C.c_float * 3),
        ("scene_hi", C.c_float * 3),
    ]


class GrtCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "rays", "segments", "rounds", "hits_gathered", "hits_shaded", "warp_nodes", "warp_leaf_prims",
        "warp_rounds", "stack_overflow", "kernel_launches", "warp_prims_hit", "warp_consume_iters", "lanes_repass",
        "order_violations", "warp_dfs_fallbacks", "warp_candidate_iters")] + [("dbg", C.c_uint64 * 8), ("warp_per_ray_passes", C.c_uint64),
                                                       ("warp_composite_rounds", C.c_uint64), ("warp_loop_iters", C.c_uint64),
                                                       ("warp_forced_flushes", C.c_uint64), ("warp_dfs_iters", C.c_uint64), ("warp_list_full", C.c_uint64)]


# every symbol include/grt.h declares (tests check that the .so exports all of them)
GRT_SYMBOLS = [
    "grtCreate", "grtDestroy", "grtGetLastError", "grtSetParticles", "grtBuildParticleAccel", "grtSetMeshes",
    "grtBuildMeshAccel", "grtLaunch", "grtLaunchTile", "grtLaunchBatch", "grtRenderHost", "grtSetDebugOutputs",
    "grtEnableCounters", "grtGetCounters", "grtGetBuildInfo", "grtValidateAccel", "grtLastLaunchMs", "grtMalloc",
    "grtFree", "grtMallocHost", "grtFreeHost", "grtMemcpyDtoH", "grtMemcpyHtoD", "grtMemset", "grtStreamCreate",
    "grtStreamDestroy", "grtStreamSynchronize", "grtDebugRadixSort", "grtSortPairsNoAlloc", "grtSetWavefront", "grtSetParticlesRaw", "grtLoadPly", "grtGetParticles",
    "grtLaunchRows", "grtGetAbiLayout", "grtCommGetUniqueId", "grtCommInit", "grtCommDestroy", "grtGatherFrames", "grtGatherBands",
]

_lib = None


def load_libgrt():
    """Load csrc/libgrt.so.  Fails loudly if the CUDA extension has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("GRT_LIBGRT", LIBGRT_PATH)  # tuning builds (tools/build_variants.sh); same ABI, still CUDA-only
    if not os.path.exists(path):
        raise GrtError(f"{path} is missing: run `make -C {_HERE}` (or __graft_entry__.build()); there is no fallback path")
    lib = C.CDLL(path)
    vp, u32, u64, i32, f32 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int, C.c_float
    sig = {
        "grtCreate": [i32, C.POINTER(vp)],
        "grtDestroy": [vp],
        "grtSetParticles": [vp, vp, u64, f32],
        "grtBuildParticleAccel": [vp, vp, C.POINTER(u64), C.POINTER(vp)],
        "grtSetMeshes": [vp, C.POINTER(GrtMeshDesc), u32],
        "grtBuildMeshAccel": [vp, vp, C.POINTER(u64), C.POINTER(vp)],
        "grtLaunch": [vp, vp, C.POINTER(GrtParams)],
        "grtLaunchTile": [vp, vp, C.POINTER(GrtParams), u32, u32, u32, u32],
        "grtLaunchBatch": [vp, vp, C.POINTER(GrtParams), u32],
        "grtRenderHost": [vp, C.POINTER(GrtParams), vp, u32, u32],
        "grtSetDebugOutputs": [vp, vp, vp, vp, u32, u32, u32, u32, u32],
        "grtEnableCounters": [vp, i32],
        "grtSetWavefront": [vp, i32],
        "grtGetCounters": [vp, C.POINTER(GrtCounters), i32],
        "grtGetBuildInfo": [vp, C.POINTER(GrtBuildInfo)],
        "grtValidateAccel": [vp, i32],
        "grtLastLaunchMs": [vp, C.POINTER(f32)],
        "grtMalloc": [vp, C.c_size_t, C.POINTER(vp)],
        "grtFree": [vp, vp],
        "grtMallocHost": [vp, C.c_size_t, C.POINTER(vp)],
        "grtFreeHost": [vp, vp],
        "grtMemcpyDtoH": [vp, vp, vp, C.c_size_t, vp],
        "grtMemcpyHtoD": [vp, vp, vp, C.c_size_t, vp],
        "grtMemset": [vp, vp, i32, C.c_size_t, vp],
        "grtStreamCreate": [vp, C.POINTER(vp)],
        "grtStreamDestroy": [vp, vp],
        "grtStreamSynchronize": [vp, vp],
        "grtDebugRadixSort": [vp, vp, vp, u64],
        "grtSortPairsNoAlloc": [vp, vp, vp, u64, vp],
        "grtSetParticlesRaw": [vp, vp, u64, u32, C.POINTER(u32), f32],
        "grtLoadPly": [vp, C.c_char_p, f32, C.POINTER(u64), C.POINTER(f32)],
        "grtGetParticles": [vp, vp, u64],
        "grtLaunchRows": [vp, vp, C.POINTER(GrtParams), u32, u32, u32],
        "grtGetAbiLayout": [C.POINTER(u64)],
        "grtCommGetUniqueId": [vp],
        "grtCommInit": [vp, i32, i32, vp],
        "grtCommDestroy": [vp],
        "grtGatherFrames": [vp, vp, vp, C.c_size_t, vp, i32],
        "grtGatherBands": [vp, vp, vp, u32, u32, u32, i32],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    lib.grtGetLastError.argtypes = [vp]
    lib.grtGetLastError.restype = C.c_char_p
    _lib = lib
    return lib


def _np_ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One grtContext (one GPU).  Thin, 1:1 over the C ABI."""

    def __init__(self, device: int = 0):
        self.lib = load_libgrt()
        h = C.c_void_p()
        rc = self.lib.grtCreate(device, C.byref(h))
        if rc:
            raise GrtError(self.lib.grtGetLastError(None).decode())
        self.h = h
        self._owned = []

    def _ck(self, rc):
        if rc:
            raise GrtError(self.lib.grtGetLastError(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            for p in self._owned:
                self.lib.grtFree(self.h, p)
            self._owned = []
            self.lib.grtDestroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- memory
    def malloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self._ck(self.lib.grtMalloc(self.h, nbytes, C.byref(p)))
        self._owned.append(p.value)
        return p.value

    def free(self, p: int):
        self._owned.remove(p)
        self._ck(self.lib.grtFree(self.h, p))

    def memset(self, p: int, value: int, nbytes: int):
        self._ck(self.lib.grtMemset(self.h, p, value, nbytes, None))

    def to_host(self, p: int, shape, dtype) -> np.ndarray:
        out = np.empty(shape, dtype=dtype)
        self._ck(self.lib.grtMemcpyDtoH(self.h, _np_ptr(out), p, out.nbytes, None))
        return out

    def to_device(self, p: int, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        self._ck(self.lib.grtMemcpyHtoD(self.h, p, _np_ptr(arr), arr.nbytes, None))

    def sync(self, stream=None):
        self._ck(self.lib.grtStreamSynchronize(self.h, stream))

    # ---- scene
    def set_particles(self, particles: np.ndarray, alpha_min: float = 0.01):
        """particles: (n, 59) float32 = n reference GaussianParticle records (236 B each)."""
        particles = np.ascontiguousarray(particles, dtype=np.float32)
        if particles.size and (particles.ndim != 2 or particles.shape[1] != PARTICLE_FLOATS):
            raise GrtError("particles must be (n, 59) float32")
        n = particles.shape[0] if particles.size else 0
        self._ck(self.lib.grtSetParticles(self.h, _np_ptr(particles) if n else None, n, alpha_min))

    def set_particles_raw(self, rows: np.ndarray, prop_offsets, alpha_min: float = 0.01):
        """rows: (n, stride_bytes) uint8 or (n, k) float32 raw PLY vertex rows; prop_offsets: 59 byte offsets
        (x y z scale_0-2 rot_0-3 opacity f_dc_0-2 f_rest_0-44).  Activation runs on the device (csrc/ingest.cu)."""
        rows = np.ascontiguousarray(rows)
        n = rows.shape[0]
        stride = rows.strides[0] if n else 236
        off = (C.c_uint32 * 59)(*[int(o) for o in prop_offsets])
        self._ck(self.lib.grtSetParticlesRaw(self.h, _np_ptr(rows) if n else None, n, stride, off, alpha_min))

    def load_ply(self, path: str, alpha_min: float = 0.01):
        """GaussianData(path) on the device: returns (vertex count, centre as float32[3])."""
        n, cen = C.c_uint64(), (C.c_float * 3)()
        self._ck(self.lib.grtLoadPly(self.h, path.encode(), alpha_min, C.byref(n), cen))
        return int(n.value), np.array(list(cen), np.float32)

    def get_particles(self, n: int) -> np.ndarray:
        out = np.empty((n, PARTICLE_FLOATS), np.float32)
        self._ck(self.lib.grtGetParticles(self.h, _np_ptr(out) if n else None, n))
        return out

    def build_particle_accel(self, stream=None):
        handle, dpart = C.c_uint64(), C.c_void_p()
        self._ck(self.lib.grtBuildParticleAccel(self.h, stream, C.byref(handle), C.byref(dpart)))
        return handle.value, dpart.value

    def set_meshes(self, meshes):
        """meshes: list of dicts(vertices (nv,3), normals (nv,3), indices (ni,), transform (4,4) glm column-major
        i.e. transform[c][r])."""
        descs = (GrtMeshDesc * max(1, len(meshes)))()
        keep = []
        for i, m in enumerate(meshes):
            v = np.ascontiguousarray(m["vertices"], dtype=np.float32).reshape(-1, 3)
            nrm = np.ascontiguousarray(m["normals"], dtype=np.float32).reshape(-1, 3)
            idx = np.ascontiguousarray(m["indices"], dtype=np.uint32).reshape(-1)
            t = np.ascontiguousarray(m["transform"], dtype=np.float32).reshape(16)
            keep += [v, nrm, idx]
            descs[i].vertices = v.ctypes.data_as(C.POINTER(C.c_float))
            descs[i].normals = nrm.ctypes.data_as(C.POINTER(C.c_float))
            descs[i].indices = idx.ctypes.data_as(C.POINTER(C.c_uint32))
            descs[i].vertex_count = v.shape[0]
            descs[i].index_count = idx.shape[0]
            for k in range(16):
                descs[i].transform[k] = float(t[k])
        self._ck(self.lib.grtSetMeshes(self.h, descs if meshes else None, len(meshes)))

    def build_mesh_accel(self, stream=None):
        handle, dm = C.c_uint64(), C.c_void_p()
        self._ck(self.lib.grtBuildMeshAccel(self.h, stream, C.byref(handle), C.byref(dm)))
        return handle.value, dm.value

    # ---- render
    def launch(self, params: GrtParams, stream=None):
        self._ck(self.lib.grtLaunch(self.h, stream, C.byref(params)))

    def launch_tile(self, params: GrtParams, x0, y0, w, h, stream=None):
        self._ck(self.lib.grtLaunchTile(self.h, stream, C.byref(params), x0, y0, w, h))

    def launch_rows(self, params: GrtParams, band_rows: int, band_stride: int, band_offset: int, stream=None):
        """One launch for the row bands b with b % band_stride == band_offset (bands of band_rows rows): rank r of G GPUs."""
        self._ck(self.lib.grtLaunchRows(self.h, stream, C.byref(params), band_rows, band_stride, band_offset))

    # ---- multi-GPU exchange (one context = one rank)
    @staticmethod
    def comm_unique_id() -> bytes:
        buf = (C.c_uint8 * 128)()
        if load_libgrt().grtCommGetUniqueId(buf):
            raise GrtError(load_libgrt().grtGetLastError(None).decode())
        return bytes(buf)

    def comm_init(self, nranks: int, rank: int, unique_id: bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._ck(self.lib.grtCommInit(self.h, nranks, rank, buf))

    def comm_destroy(self):
        self._ck(self.lib.grtCommDestroy(self.h))

    def gather_frames(self, d_send: int, nbytes: int, d_recv_all, root: int = 0, stream=None):
        self._ck(self.lib.grtGatherFrames(self.h, stream, d_send, nbytes, d_recv_all, root))

    def gather_bands(self, d_frame: int, width: int, height: int, band_rows: int = 8, root: int = 0, stream=None):
        self._ck(self.lib.grtGatherBands(self.h, stream, d_frame, width, height, band_rows, root))

    def launch_batch(self, params_list, stream=None):
        arr = (GrtParams * len(params_list))(*params_list)
        self._ck(self.lib.grtLaunchBatch(self.h, stream, arr, len(params_list)))

    def render_host(self, params: GrtParams, out: np.ndarray, y0: int = 0, h: int = 0):
        """out: (height, width, 3) uint8 host array (pinned or pageable); rows [y0, y0+h) are filled."""
        assert out.dtype == np.uint8 and out.flags["C_CONTIGUOUS"]
        self._ck(self.lib.grtRenderHost(self.h, C.byref(params), _np_ptr(out), y0, h))

    def pinned_empty(self, shape, dtype=np.uint8) -> np.ndarray:
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        self._ck(self.lib.grtMallocHost(self.h, n, C.byref(p)))
        buf = (C.c_uint8 * n).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    # ---- side channels
    def set_debug_outputs(self, d_rgb32f=None, d_trace_ids=None, d_trace_counts=None, rect=(0, 0, 0, 0), trace_max=0):
        self._ck(self.lib.grtSetDebugOutputs(self.h, d_rgb32f, d_trace_ids, d_trace_counts, *rect, trace_max))

    def set_wavefront(self, on: bool):
        self._ck(self.lib.grtSetWavefront(self.h, int(on)))

    def enable_counters(self, on: bool):
        self._ck(self.lib.grtEnableCounters(self.h, int(on)))

    def counters(self, reset=False) -> dict:
        c = GrtCounters()
        self._ck(self.lib.grtGetCounters(self.h, C.byref(c), int(reset)))
        d = {n: int(getattr(c, n)) for n, _ in GrtCounters._fields_ if n != "dbg"}
        d["dbg"] = [int(v) for v in c.dbg]
        return d

    def build_info(self) -> dict:
        b = GrtBuildInfo()
        self._ck(self.lib.grtGetBuildInfo(self.h, C.byref(b)))
        d = {n: getattr(b, n) for n, _ in GrtBuildInfo._fields_}
        d["scene_lo"] = list(b.scene_lo)
        d["scene_hi"] = list(b.scene_hi)
        return d

    def validate_accel(self, which=0):
        self._ck(self.lib.grtValidateAccel(self.h, which))

    def last_launch_ms(self) -> float:
        ms = C.c_float()
        self._ck(self.lib.grtLastLaunchMs(self.h, C.byref(ms)))
        return float(ms.value)

    def radix_sort(self, keys: np.ndarray, vals: np.ndarray):
        keys = np.ascontiguousarray(keys, dtype=np.uint64).copy()
        vals = np.ascontiguousarray(vals, dtype=np.uint32).copy()
        self._ck(self.lib.grtDebugRadixSort(self.h, _np_ptr(keys), _np_ptr(vals), keys.shape[0]))
        return keys, vals

    def sort_pairs_noalloc(self, d_keys: int, d_vals: int, n: int, stream=None):
        """In-place device sort of n (uint64 key, uint32 value) pairs through the allocation-free entry point."""
        self._ck(self.lib.grtSortPairsNoAlloc(self.h, d_keys, d_vals, n, stream))


from .scene import (  # noqa: E402
    synth_ply_properties, activate, write_ply, read_ply_properties, PLY_PROPERTIES, make_params, camera_uvw,
    default_camera, create_plane, create_sphere, translate, rotate_y, scale_mat,
)
from .tracer import GaussianTracer, CUDAOutputBuffer, Camera  # noqa: E402
