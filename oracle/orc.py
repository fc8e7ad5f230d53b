"""ctypes wrapper around oracle/liboracle.so (and oracle/_ref/libref_host.so when present).

TEST INFRASTRUCTURE ONLY: import this from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs -- never from the product package.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
REF_PATH = os.path.join(_HERE, "_ref", "libref_host.so")

FLAG_TIE, FLAG_ALPHA, FLAG_TRANSMIT, FLAG_MESH = 1, 2, 4, 8


class OrcParams(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("sh_degree_max", C.c_uint32),
                ("eye", C.c_float * 3), ("U", C.c_float * 3), ("V", C.c_float * 3), ("W", C.c_float * 3),
                ("t_min", C.c_float), ("t_max", C.c_float), ("minTransmittance", C.c_float), ("alpha_min", C.c_float),
                ("mode_fisheye", C.c_int32), ("type", C.c_int32)]


class OrcMesh(C.Structure):
    _fields_ = [("vertices", C.POINTER(C.c_float)), ("normals", C.POINTER(C.c_float)), ("faces", C.POINTER(C.c_uint32)),
                ("n_vertices", C.c_uint32), ("n_faces", C.c_uint32)]


class OrcStats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("segments", C.c_uint64), ("rounds", C.c_uint64), ("hits_gathered", C.c_uint64),
                ("hits_shaded", C.c_uint64), ("flagged", C.c_uint64), ("nodes_visited", C.c_uint64),
                ("proxies_tested", C.c_uint64), ("seconds", C.c_double), ("threads", C.c_uint32)]


class OrcRenderArgs(C.Structure):
    _fields_ = [("params", OrcParams), ("K", C.c_uint32), ("brute_force", C.c_int32), ("use_triangles", C.c_int32),
                ("threads", C.c_uint32), ("x0", C.c_uint32), ("y0", C.c_uint32), ("w", C.c_uint32), ("h", C.c_uint32),
                ("out_rgb8", C.c_void_p), ("out_rgb32f", C.c_void_p), ("out_flags", C.c_void_p), ("out_hg", C.c_void_p),
                ("trace_pixels", C.c_void_p), ("n_trace_pixels", C.c_uint32), ("trace_max", C.c_uint32),
                ("trace_ids", C.c_void_p), ("trace_counts", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} missing: run `make -C {_HERE}`")
        L = C.CDLL(LIB_PATH)
        L.orc_scene_create.argtypes = [C.c_void_p, C.c_uint64, C.c_float]
        L.orc_scene_create.restype = C.c_void_p
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_scene_set_meshes.argtypes = [C.c_void_p, C.POINTER(OrcMesh), C.c_uint32]
        L.orc_scene_live_count.argtypes = [C.c_void_p]
        L.orc_scene_live_count.restype = C.c_uint64
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(OrcRenderArgs), C.POINTER(OrcStats)]
        L.orc_render.restype = C.c_int
        L.orc_compute_response.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_compute_response.restype = C.c_float
        L.orc_sh_radiance.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]
        L.orc_quantize.argtypes = [C.c_float]
        L.orc_quantize.restype = C.c_uint8
        L.orc_get_ray.argtypes = [C.POINTER(OrcParams), C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_get_ray.restype = C.c_int
        L.orc_proxy_hit.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_proxy_hit.restype = C.c_int
        L.orc_icosahedron.argtypes = [C.c_void_p, C.c_void_p]
        L.orc_refract.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_uint32)]
        L.orc_transform_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def params_from_grt(gp) -> OrcParams:
    """Copy the numeric fields of a GrtParams (reference Params layout) into OrcParams."""
    p = OrcParams()
    p.width, p.height, p.sh_degree_max = gp.width, gp.height, gp.sh_degree_max
    for k in range(3):
        p.eye[k], p.U[k], p.V[k], p.W[k] = gp.eye[k], gp.U[k], gp.V[k], gp.W[k]
    p.t_min, p.t_max, p.minTransmittance, p.alpha_min = gp.t_min, gp.t_max, gp.minTransmittance, gp.alpha_min
    p.mode_fisheye = int(gp.mode_fisheye)
    p.type = int(gp.type)
    return p


class Scene:
    def __init__(self, particles: np.ndarray, alpha_min: float = 0.01):
        self.particles = np.ascontiguousarray(particles, dtype=np.float32)
        assert self.particles.ndim == 2 and self.particles.shape[1] == 59
        self.h = lib().orc_scene_create(_p(self.particles), self.particles.shape[0], alpha_min)
        self._mesh_keep = []

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_scene_destroy(self.h)
            self.h = None

    def live_count(self):
        return int(lib().orc_scene_live_count(self.h))

    def set_meshes(self, meshes):
        """meshes: same dicts as Context.set_meshes (object-space verts/normals, indices, glm transform)."""
        arr = (OrcMesh * max(1, len(meshes)))()
        keep = []
        for i, m in enumerate(meshes):
            v = np.ascontiguousarray(m["vertices"], np.float32).reshape(-1, 3)
            n = np.ascontiguousarray(m["normals"], np.float32).reshape(-1, 3)
            f = np.ascontiguousarray(m["indices"], np.uint32).reshape(-1)
            t = np.ascontiguousarray(m["transform"], np.float32).reshape(16)
            vw, nw = np.empty_like(v), np.empty_like(n)
            lib().orc_transform_mesh(_p(v), _p(n), v.shape[0], _p(t), _p(vw), _p(nw))
            keep += [vw, nw, f]
            arr[i].vertices = vw.ctypes.data_as(C.POINTER(C.c_float))
            arr[i].normals = nw.ctypes.data_as(C.POINTER(C.c_float))
            arr[i].faces = f.ctypes.data_as(C.POINTER(C.c_uint32))
            arr[i].n_vertices = v.shape[0]
            arr[i].n_faces = f.shape[0] // 3
        lib().orc_scene_set_meshes(self.h, arr, len(meshes))
        self._mesh_keep = keep

    def render(self, params: OrcParams, *, K=7, brute_force=False, use_triangles=False, threads=0, rect=None,
               want_float=True, want_flags=True, want_hg=False, trace_pixels=None, trace_max=0):
        W, H = params.width, params.height
        a = OrcRenderArgs()
        a.params = params
        a.K, a.brute_force, a.use_triangles, a.threads = K, int(brute_force), int(use_triangles), threads
        if rect:
            a.x0, a.y0, a.w, a.h = rect
        out = {}
        rgb8 = np.zeros((H, W, 3), np.uint8)
        a.out_rgb8 = rgb8.ctypes.data
        out["rgb8"] = rgb8
        if want_float:
            f = np.zeros((H, W, 3), np.float32)
            a.out_rgb32f = f.ctypes.data
            out["rgb32f"] = f
        if want_flags:
            fl = np.zeros((H, W), np.uint8)
            a.out_flags = fl.ctypes.data
            out["flags"] = fl
        if want_hg:
            hg = np.zeros((H, W), np.uint16)
            a.out_hg = hg.ctypes.data
            out["hg"] = hg
        if trace_pixels is not None and trace_max:
            tp = np.ascontiguousarray(trace_pixels, np.uint32)
            ids = np.full((tp.shape[0], trace_max), 0xFFFFFFFF, np.uint32)
            cnt = np.zeros(tp.shape[0], np.uint32)
            a.trace_pixels, a.n_trace_pixels, a.trace_max = tp.ctypes.data, tp.shape[0], trace_max
            a.trace_ids, a.trace_counts = ids.ctypes.data, cnt.ctypes.data
            out["trace_ids"], out["trace_counts"] = ids, cnt
            out["_tp"] = tp
        st = OrcStats()
        rc = lib().orc_render(self.h, C.byref(a), C.byref(st))
        if rc:
            raise RuntimeError(f"orc_render failed rc={rc}")
        out["stats"] = {n: getattr(st, n) for n, _ in OrcStats._fields_}
        return out


# ---- KAT hooks
def compute_response(rec59, o, d):
    r = np.ascontiguousarray(rec59, np.float32)
    o = np.ascontiguousarray(o, np.float32)
    d = np.ascontiguousarray(d, np.float32)
    return float(lib().orc_compute_response(_p(r), _p(o), _p(d)))


def sh_radiance(rec59, d, deg):
    r = np.ascontiguousarray(rec59, np.float32)
    d = np.ascontiguousarray(d, np.float32)
    out = np.zeros(3, np.float32)
    lib().orc_sh_radiance(_p(r), _p(d), deg, _p(out))
    return out


def quantize(x):
    return int(lib().orc_quantize(float(x)))


def get_ray(params: OrcParams, x, y):
    o, d = np.zeros(3, np.float32), np.zeros(3, np.float32)
    ok = lib().orc_get_ray(C.byref(params), x, y, _p(o), _p(d))
    return bool(ok), o, d


def proxy_hit(rec59, alpha_min, o, d, use_triangles=False):
    r = np.ascontiguousarray(rec59, np.float32)
    o = np.ascontiguousarray(o, np.float32)
    d = np.ascontiguousarray(d, np.float32)
    t = np.zeros(2, np.float32)
    n = lib().orc_proxy_hit(_p(r), alpha_min, _p(o), _p(d), int(use_triangles), _p(t))
    return n, t


def icosahedron():
    v, i = np.zeros((12, 3), np.float32), np.zeros(60, np.uint32)
    lib().orc_icosahedron(_p(v), _p(i))
    return v, i


def refract(d, n, eta, t_hit=1.0, bounces=0):
    d = np.ascontiguousarray(d, np.float32)
    n = np.ascontiguousarray(n, np.float32)
    out = np.zeros(3, np.float32)
    t, b = C.c_float(t_hit), C.c_uint32(bounces)
    lib().orc_refract(_p(d), _p(n), eta, _p(out), C.byref(t), C.byref(b))
    return out, float(t.value), int(b.value)


# ---- the reference's own host code (oracle/_ref), when built
def ref_available():
    return os.path.exists(REF_PATH)


def ref_load_ply(path):
    L = C.CDLL(REF_PATH)
    L.ref_load_ply.argtypes = [C.c_char_p, C.POINTER(C.c_ulonglong), C.c_void_p]
    L.ref_load_ply.restype = C.c_void_p
    L.ref_free.argtypes = [C.c_void_p]
    n = C.c_ulonglong()
    center = np.zeros(3, np.float32)
    p = L.ref_load_ply(path.encode(), C.byref(n), _p(center))
    if not p:
        raise RuntimeError("reference GaussianData threw")
    arr = np.frombuffer((C.c_uint8 * (n.value * 236)).from_address(p), dtype=np.float32).reshape(n.value, 59).copy()
    L.ref_free(p)
    return arr, center


def ref_camera_uvw(eye, lookat, up, fovy, aspect):
    L = C.CDLL(REF_PATH)
    L.ref_camera_uvw.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    e, l, u = (np.ascontiguousarray(x, np.float32) for x in (eye, lookat, up))
    U, V, W = np.zeros(3, np.float32), np.zeros(3, np.float32), np.zeros(3, np.float32)
    L.ref_camera_uvw(_p(e), _p(l), _p(u), fovy, aspect, _p(U), _p(V), _p(W))
    return U, V, W


# ---- batch forms of the KAT hooks (oracle side) -------------------------------------------------------------------
def _f32c(a):
    return np.ascontiguousarray(a, np.float32)


def _u32c(a):
    return np.ascontiguousarray(a, np.uint32)


def compute_response_batch(particles, pidx, o, d):
    ps, pidx, o, d = _f32c(particles), _u32c(pidx), _f32c(o), _f32c(d)
    out = np.zeros(len(pidx), np.float32)
    f = lib().orc_compute_response_batch
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = None
    f(_p(ps), len(pidx), _p(pidx), _p(o), _p(d), _p(out))
    return out


def sh_radiance_batch(particles, pidx, d, deg, normalise=True):
    ps, pidx, d = _f32c(particles), _u32c(pidx), _f32c(d)
    out = np.zeros((len(pidx), 3), np.float32)
    f = lib().orc_sh_radiance_batch
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
    f.restype = None
    f(_p(ps), len(pidx), _p(pidx), _p(d), deg, int(normalise), _p(out))
    return out


def quantize_batch(x):
    x = _f32c(x)
    out = np.zeros(x.shape, np.uint8)
    f = lib().orc_quantize_batch
    f.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p]
    f.restype = None
    f(x.size, _p(x), _p(out))
    return out


def get_rays(params: OrcParams, xs, ys):
    xs, ys = _u32c(xs), _u32c(ys)
    o, d = np.zeros((len(xs), 3), np.float32), np.zeros((len(xs), 3), np.float32)
    valid = np.zeros(len(xs), np.uint8)
    f = lib().orc_get_rays
    f.argtypes = [C.POINTER(OrcParams), C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = None
    f(C.byref(params), len(xs), _p(xs), _p(ys), _p(o), _p(d), _p(valid))
    return o, d, valid.astype(bool)


def refract_batch(d, n, eta, t_hit, bounces):
    d, n = _f32c(d), _f32c(n)
    out = np.zeros_like(d)
    t, b = _f32c(t_hit).copy(), _u32c(bounces).copy()
    f = lib().orc_refract_batch
    f.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = None
    f(len(d), _p(d), _p(n), eta, _p(out), _p(t), _p(b))
    return out, t, b


def mirror_batch(d, n, bounces):
    d, n = _f32c(d), _f32c(n)
    out = np.zeros_like(d)
    b = _u32c(bounces).copy()
    f = lib().orc_mirror_batch
    f.argtypes = [C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = None
    f(len(d), _p(d), _p(n), _p(out), _p(b))
    return out, b


def proxy_data(particles, alpha_min=0.01):
    ps = _f32c(particles)
    n = ps.shape[0]
    A, s = np.zeros((n, 3, 3), np.float32), np.zeros(n, np.float32)
    wv, box, live = np.zeros((n, 12, 3), np.float32), np.zeros((n, 6), np.float32), np.zeros(n, np.uint8)
    f = lib().orc_proxy_data
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_float] + [C.c_void_p] * 5
    f.restype = None
    f(_p(ps), n, alpha_min, _p(A), _p(s), _p(wv), _p(box), _p(live))
    return dict(A=A, s=s, wv=wv, box=box, live=live.astype(bool))


def proxy_hit_batch(particles, pidx, o, d, alpha_min=0.01):
    ps, pidx, o, d = _f32c(particles), _u32c(pidx), _f32c(o), _f32c(d)
    t2, hit = np.zeros((len(pidx), 2), np.float32), np.zeros(len(pidx), np.uint8)
    f = lib().orc_proxy_hit_batch
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    f.restype = None
    f(_p(ps), len(pidx), _p(pidx), alpha_min, _p(o), _p(d), _p(t2), _p(hit))
    return t2, hit.astype(bool)


def soup_hits(verts, idx60, o, d):
    verts, idx60, o, d = _f32c(verts), _u32c(idx60), _f32c(o), _f32c(d)
    n = verts.shape[0]
    t2, hit = np.zeros((n, 2), np.float32), np.zeros(n, np.uint8)
    f = lib().orc_soup_hits
    f.argtypes = [C.c_uint64] + [C.c_void_p] * 6
    f.restype = None
    f(n, _p(verts), _p(idx60), _p(o), _p(d), _p(t2), _p(hit))
    return t2, hit.astype(bool)


# ---- the reference's own hot-path math (oracle/_ref, ref_math_shim.cpp) ---------------------------------------------
_ref = None


def ref():
    """ctypes handle of oracle/_ref/libref_host.so with the math entry points typed (raises if the library is absent or older
    than ref_math_shim.cpp: rebuild with `make -C oracle ref` in a container that has /root/reference)."""
    global _ref
    if _ref is None:
        L = C.CDLL(REF_PATH)
        if not hasattr(L, "ref_layout"):
            raise RuntimeError(f"{REF_PATH} predates ref_math_shim.cpp: run `make -C {_HERE} ref`")
        V = C.c_void_p
        L.ref_layout.argtypes = [V]
        L.ref_constants.argtypes = [V]
        L.ref_param_defaults.argtypes = [C.c_float, V, V]
        L.ref_quantize.argtypes = [C.c_uint64, V, V]
        L.ref_get_rays.argtypes = [C.c_int, C.c_uint64, V, V, V, V, V, V, C.c_int, C.c_int, V, V, V]
        L.ref_compute_response.argtypes = [V, C.c_uint64, V, V, V, V]
        L.ref_sh_radiance.argtypes = [V, C.c_uint64, V, V, C.c_uint32, C.c_int, V]
        L.ref_refract.argtypes = [C.c_uint64, V, V, C.c_float, V, V, V]
        L.ref_mirror.argtypes = [C.c_uint64, V, V, V, V]
        L.ref_proxy_transforms.argtypes = [V, C.c_uint64, C.c_float, V]
        L.ref_icosahedron.argtypes = [V, V]
        L.ref_primitive.argtypes = [C.c_int, V, V, V, V, V, V]
        L.ref_normal_transform.argtypes = [C.c_uint64, V, V, V]
        L.ref_pack_transform.argtypes = [V, V]
        L.ref_placement.argtypes = [V, V, V]
        L.ref_trs.argtypes = [V, C.c_float, C.c_float, V]
        for n in ("ref_layout ref_constants ref_param_defaults ref_quantize ref_get_rays ref_compute_response ref_sh_radiance ref_refract "
                  "ref_mirror ref_proxy_transforms ref_icosahedron ref_primitive ref_normal_transform ref_pack_transform ref_placement ref_trs").split():
            getattr(L, n).restype = None
        _ref = L
    return _ref


REF_LAYOUT_NAMES = ["sizeof_Params", "output_buffer", "width", "height", "sh_degree_max", "eye", "U", "V", "W", "t_min", "t_max",
                    "minTransmittance", "alpha_min", "handle", "d_particles", "mesh_handle", "mode_fisheye", "d_meshes", "type", "traceState",
                    "sizeof_GaussianParticle", "gp_position", "gp_scale", "gp_rotation", "gp_opacity", "gp_sh",
                    "sizeof_Mesh", "mesh_faces", "mesh_vertex_normals", "mesh_type_enum", "MaxNumHitPerTrace", "timeout_bounces"]


def ref_layout():
    out = np.zeros(32, np.uint64)
    ref().ref_layout(_p(out))
    return dict(zip(REF_LAYOUT_NAMES, (int(v) for v in out)))


def ref_constants():
    out = np.zeros(24, np.float32)
    ref().ref_constants(_p(out))
    return out


def ref_param_defaults(alpha_min=0.01):
    f, i = np.zeros(4, np.float32), np.zeros(3, np.int32)
    ref().ref_param_defaults(alpha_min, _p(f), _p(i))
    return f, i


def ref_quantize(x):
    x = _f32c(x)
    out = np.zeros(x.shape, np.uint8)
    ref().ref_quantize(x.size, _p(x), _p(out))
    return out


def ref_get_rays(fisheye, xs, ys, U, V, W, eye, width, height):
    xs, ys = _u32c(xs), _u32c(ys)
    U, V, W, eye = _f32c(U), _f32c(V), _f32c(W), _f32c(eye)
    o, d = np.zeros((len(xs), 3), np.float32), np.zeros((len(xs), 3), np.float32)
    valid = np.zeros(len(xs), np.uint8)
    ref().ref_get_rays(int(fisheye), len(xs), _p(xs), _p(ys), _p(U), _p(V), _p(W), _p(eye), width, height, _p(o), _p(d), _p(valid))
    return o, d, valid.astype(bool)


def ref_compute_response(particles, pidx, o, d):
    ps, pidx, o, d = _f32c(particles), _u32c(pidx), _f32c(o), _f32c(d)
    out = np.zeros(len(pidx), np.float32)
    ref().ref_compute_response(_p(ps), len(pidx), _p(pidx), _p(o), _p(d), _p(out))
    return out


def ref_sh_radiance(particles, pidx, d, deg, normalise=True):
    ps, pidx, d = _f32c(particles), _u32c(pidx), _f32c(d)
    out = np.zeros((len(pidx), 3), np.float32)
    ref().ref_sh_radiance(_p(ps), len(pidx), _p(pidx), _p(d), deg, int(normalise), _p(out))
    return out


def ref_refract(d, n, eta, t_hit, bounces):
    d, n = _f32c(d), _f32c(n)
    out = np.zeros_like(d)
    t, b = _f32c(t_hit).copy(), _u32c(bounces).copy()
    ref().ref_refract(len(d), _p(d), _p(n), eta, _p(out), _p(t), _p(b))
    return out, t, b


def ref_mirror(d, n, bounces):
    d, n = _f32c(d), _f32c(n)
    out = np.zeros_like(d)
    b = _u32c(bounces).copy()
    ref().ref_mirror(len(d), _p(d), _p(n), _p(out), _p(b))
    return out, b


def ref_proxy_transforms(particles, alpha_min=0.01):
    ps = _f32c(particles)
    out = np.zeros((ps.shape[0], 3, 4), np.float32)
    ref().ref_proxy_transforms(_p(ps), ps.shape[0], alpha_min, _p(out))
    return out


def ref_icosahedron():
    v, i = np.zeros((12, 3), np.float32), np.zeros(60, np.uint32)
    ref().ref_icosahedron(_p(v), _p(i))
    return v, i


def ref_primitive(kind, position):
    pos = _f32c(position)
    counts, T = np.zeros(2, np.uint64), np.zeros(16, np.float32)
    ref().ref_primitive(kind, _p(pos), None, None, None, _p(counts), _p(T))
    v, n = np.zeros((int(counts[0]), 3), np.float32), np.zeros((int(counts[0]), 3), np.float32)
    idx = np.zeros(int(counts[1]), np.uint32)
    ref().ref_primitive(kind, _p(pos), _p(v), _p(n), _p(idx), _p(counts), _p(T))
    return dict(vertices=v, normals=n, indices=idx, transform=T.reshape(4, 4))


def ref_normal_transform(normals, transform_colmajor16):
    nn, T = _f32c(normals), _f32c(transform_colmajor16).reshape(16)
    out = np.zeros_like(nn)
    ref().ref_normal_transform(nn.shape[0], _p(nn), _p(T), _p(out))
    return out


def ref_pack_transform(transform_colmajor16):
    T, out = _f32c(transform_colmajor16).reshape(16), np.zeros((3, 4), np.float32)
    ref().ref_pack_transform(_p(T), _p(out))
    return out


def ref_placement(eye, lookat):
    e, l, out = _f32c(eye), _f32c(lookat), np.zeros(3, np.float32)
    ref().ref_placement(_p(e), _p(l), _p(out))
    return out


def ref_trs(t, rot_y_deg, s):
    t, out = _f32c(t), np.zeros(16, np.float32)
    ref().ref_trs(_p(t), rot_y_deg, s, _p(out))
    return out.reshape(4, 4)
