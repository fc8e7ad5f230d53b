"""Test/bench harness helpers: synthetic scene generator, PLY I/O, activation, camera, primitives.

Nothing here is on the hot path.  Each function names the reference code whose behaviour it mirrors.
The synthetic scene is the generator of SURVEY.md section 8d in counter-based form (every random draw is
``splitmix64(seed, stream, particle, slot)`` so that it vectorises in numpy and is reproducible anywhere).
"""
from __future__ import annotations

import numpy as np

# INRIA / 3DGRUT property order (62 floats, 248 B per vertex)
PLY_PROPERTIES = (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{i}" for i in range(3)] + [f"f_rest_{i}" for i in range(45)]
                  + ["opacity"] + [f"scale_{i}" for i in range(3)] + [f"rot_{i}" for i in range(4)])
_COL = {n: i for i, n in enumerate(PLY_PROPERTIES)}


def _splitmix64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15)).astype(np.uint64)
    z = x.copy()
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def _u01(seed: int, stream: int, idx: np.ndarray, slot: int) -> np.ndarray:
    with np.errstate(over="ignore"):
        key = (np.uint64(seed) * np.uint64(0xD1342543DE82EF95) + (np.uint64(stream) << np.uint64(56))
               + idx.astype(np.uint64) * np.uint64(256) + np.uint64(slot))
        h = _splitmix64(_splitmix64(key))
    return (h >> np.uint64(11)).astype(np.float64) * (2.0 ** -53)


def _gauss(seed: int, idx: np.ndarray, k: int) -> np.ndarray:
    u1 = _u01(seed, 1, idx, 2 * k)
    u2 = _u01(seed, 1, idx, 2 * k + 1)
    return np.sqrt(-2.0 * np.log(1.0 - u1)) * np.cos(2.0 * np.pi * u2)


def synth_ply_properties(n: int, seed: int) -> np.ndarray:
    """(n, 62) float32 raw PLY properties of the synthetic scene: 60 % unit-sphere shell,
    25 % ground disk, 15 % far background; flattened anisotropic splats; SH degree 3; shuffled order."""
    idx = np.arange(n, dtype=np.uint64)
    nA = int(0.60 * n)
    nB = int(0.25 * n)
    nC = n - nA - nB
    pop = np.where(idx < nA, 0, np.where(idx < nA + nB, 1, 2))
    g = lambda k: _gauss(seed, idx, k)  # noqa: E731
    u = lambda k: _u01(seed, 2, idx, k)  # noqa: E731
    out = np.zeros((n, 62), dtype=np.float64)
    # positions
    v = np.stack([g(0), g(1), g(2)], axis=1)
    v /= np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-12)
    posA = v * (1.0 + 0.03 * g(3))[:, None]
    r = 4.0 * np.sqrt(u(0))
    ang = 2.0 * np.pi * u(1)
    posB = np.stack([r * np.cos(ang), -1.0 + 0.02 * g(3), r * np.sin(ang)], axis=1)
    posC = v * (4.0 + 4.0 * u(2))[:, None]
    pos = np.where((pop == 0)[:, None], posA, np.where((pop == 1)[:, None], posB, posC))
    out[:, 0:3] = pos
    # scales
    c = 0.6
    s0 = np.array([c * np.sqrt(4 * np.pi / max(nA, 1)), c * np.sqrt(16 * np.pi / max(nB, 1)), 4 * c * np.sqrt(144 * np.pi / max(nC, 1))])
    ls = np.log(s0)[pop]
    out[:, _COL["scale_0"]] = ls + 0.4 * g(4)
    out[:, _COL["scale_1"]] = ls + 0.4 * g(5)
    out[:, _COL["scale_2"]] = ls + 0.4 * g(6) + np.log(0.25)
    for k in range(4):
        out[:, _COL["rot_0"] + k] = g(7 + k)
    out[:, _COL["opacity"]] = 0.5 + 2.0 * g(11)
    for k in range(3):
        out[:, _COL["f_dc_0"] + k] = 1.2 * g(12 + k)
    for ch in range(3):
        for k in range(15):
            amp = 0.15 if k < 3 else (0.08 if k < 8 else 0.04)
            out[:, _COL["f_rest_0"] + ch * 15 + k] = amp * g(15 + ch * 15 + k)
    # shuffle (no spatial coherence in memory, like a trained PLY)
    order = np.argsort(_u01(seed, 3, idx, 0), kind="stable")
    return out[order].astype(np.float32)


def activate(props: np.ndarray) -> np.ndarray:
    """Raw PLY properties (n,62) -> (n,59) float32 GaussianParticle records; mirrors GaussianData::parse
    (src/GaussianData.cpp:97-131): exp(scale), quaternion normalised with rot_0 = w (memory order x,y,z,w),
    sigmoid(opacity), sh[k] = (f_rest[k-1], f_rest[k+14], f_rest[k+29])."""
    p = np.ascontiguousarray(props, dtype=np.float32)
    n = p.shape[0]
    rec = np.zeros((n, 59), dtype=np.float32)
    rec[:, 0:3] = p[:, 0:3]
    rec[:, 3:6] = np.exp(p[:, _COL["scale_0"]:_COL["scale_0"] + 3])
    r = p[:, _COL["rot_0"]:_COL["rot_0"] + 4]
    norm = np.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    rec[:, 6] = r[:, 1] / norm
    rec[:, 7] = r[:, 2] / norm
    rec[:, 8] = r[:, 3] / norm
    rec[:, 9] = r[:, 0] / norm
    rec[:, 10] = np.float32(1.0) / (np.float32(1.0) + np.exp(-p[:, _COL["opacity"]]))
    rec[:, 11:14] = p[:, _COL["f_dc_0"]:_COL["f_dc_0"] + 3]
    fr = p[:, _COL["f_rest_0"]:_COL["f_rest_0"] + 45]
    for k in range(1, 16):
        rec[:, 11 + 3 * k + 0] = fr[:, k - 1]
        rec[:, 11 + 3 * k + 1] = fr[:, k + 14]
        rec[:, 11 + 3 * k + 2] = fr[:, k + 29]
    return rec


def write_ply(path: str, props: np.ndarray, names=PLY_PROPERTIES):
    """binary_little_endian PLY with float properties, readable by the reference's happly loader."""
    props = np.ascontiguousarray(props, dtype="<f4")
    with open(path, "wb") as f:
        hdr = ["ply", "format binary_little_endian 1.0", f"element vertex {props.shape[0]}"]
        hdr += [f"property float {n}" for n in names] + ["end_header"]
        f.write(("\n".join(hdr) + "\n").encode("ascii"))
        f.write(props.tobytes())


def read_ply_properties(path: str):
    """Minimal binary-LE float-only reader for the harness (the C++ host library has the full reader)."""
    with open(path, "rb") as f:
        names, n = [], 0
        while True:
            line = f.readline().decode("ascii").strip()
            if line.startswith("element vertex"):
                n = int(line.split()[-1])
            elif line.startswith("property float"):
                names.append(line.split()[-1])
            elif line == "end_header":
                break
        data = np.frombuffer(f.read(n * len(names) * 4), dtype="<f4").reshape(n, len(names))
    return names, data


# ------------------------------------------------------------------------------------------ camera
def _f32(x):
    return np.asarray(x, dtype=np.float32)


def _cross(a, b):
    return _f32([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def _normalize(v):
    inv = np.float32(1.0) / np.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2], dtype=np.float32)
    return v * inv


def camera_uvw(eye, lookat, up, fov_y_deg: float, aspect: float):
    """Camera::UVWFrame (src/Camera.cpp:3-13) in float32."""
    eye, lookat, up = _f32(eye), _f32(lookat), _f32(up)
    W = lookat - eye
    wlen = np.sqrt(W[0] * W[0] + W[1] * W[1] + W[2] * W[2], dtype=np.float32)
    U = _normalize(_cross(W, up))
    V = _normalize(_cross(U, W))
    vlen = np.float32(wlen * np.tan(np.float32(0.5) * np.float32(fov_y_deg) * np.float32(np.pi) / np.float32(180.0), dtype=np.float32))
    V = V * vlen
    U = U * np.float32(vlen * np.float32(aspect))
    return U.astype(np.float32), V.astype(np.float32), W.astype(np.float32)


def gaussian_center(particles: np.ndarray) -> np.ndarray:
    """GaussianData::getCenter (src/GaussianData.cpp:140-152): sequential float32 sums, then one division."""
    pos = np.ascontiguousarray(particles[:, 0:3], dtype=np.float32)
    sums = np.add.accumulate(pos, axis=0, dtype=np.float32)[-1]
    return (sums / np.float32(len(pos))).astype(np.float32)


def default_camera(particles: np.ndarray):
    """GUI defaults (src/gui.cpp:50-56): eye (0,0,3), lookat = mean particle position, up (0,1,0), fovY 60."""
    center = gaussian_center(particles) if len(particles) else _f32([0, 0, 0])
    return dict(eye=_f32([0, 0, 3]), lookat=center, up=_f32([0, 1, 0]), fov_y=60.0)


def make_params(width, height, eye, U, V, W, *, sh_degree_max=0, fisheye=False, mesh_type=0, handle=0, d_particles=0,
                mesh_handle=0, d_meshes=0, output=0, alpha_min=0.01):
    """Params with the reference defaults of initializeParams (src/GaussianTracer.cpp:475-486)."""
    from . import GrtParams
    p = GrtParams()
    p.output_buffer = output or None
    p.width, p.height, p.sh_degree_max = int(width), int(height), int(sh_degree_max)
    for k in range(3):
        p.eye[k], p.U[k], p.V[k], p.W[k] = float(eye[k]), float(U[k]), float(V[k]), float(W[k])
    p.t_min, p.t_max, p.minTransmittance, p.alpha_min = 1e-3, 1e5, 0.001, alpha_min
    p.handle, p.d_particles = handle, d_particles or None
    p.mesh_handle, p.d_meshes = mesh_handle, d_meshes or None
    p.mode_fisheye = 1 if fisheye else 0
    p.type = int(mesh_type)
    p.traceState = None
    return p


# ------------------------------------------------------------------------------------------ primitives / transforms
# Transforms are returned in glm memory layout: T[c][r] (column-major), so T.reshape(16) is what glm stores.
def translate(p):
    T = np.eye(4, dtype=np.float32)
    T[3, 0:3] = _f32(p)
    return T


def rotate_y(deg):
    a = np.float32(np.deg2rad(deg))
    c, s = np.cos(a, dtype=np.float32), np.sin(a, dtype=np.float32)
    M = np.eye(4, dtype=np.float32)  # math (row, col)
    M[0, 0], M[0, 2], M[2, 0], M[2, 2] = c, s, -s, c
    return M.T.copy()


def scale_mat(s):
    T = np.eye(4, dtype=np.float32)
    T[0, 0] = T[1, 1] = T[2, 2] = np.float32(s)
    return T


def compose(*glm_mats):
    """glm product A*B*C of glm-layout matrices."""
    M = np.eye(4, dtype=np.float32)
    for g in glm_mats:
        M = (M @ g.T).astype(np.float32)
    return M.T.copy()


def primitive_position(eye, lookat):
    """Where createPlane / createSphere put a new primitive: 0.75 eye + 0.25 lookat (src/GaussianTracer.cpp:580-588), in the
    reference's order of operations."""
    eye, lookat = _f32(eye), _f32(lookat)
    cw = np.float32(0.75)
    gw = np.float32(1.0) - cw
    return (lookat * gw + eye * cw).astype(np.float32)


def create_plane(position):
    """Primitives::createPlane (src/geometry/Primitives.cpp:6-61): 0.3 x 0.5 quad, normal +z, 2 triangles."""
    w, h = np.float32(0.3), np.float32(0.5)
    corner = _f32([-w * np.float32(0.5), -h * np.float32(0.5), 0.0])
    verts = []
    for j in range(2):
        for i in range(2):
            verts.append(corner + _f32([np.float32(i) * w, np.float32(j) * h, 0.0]))
    verts = np.stack(verts).astype(np.float32)
    normals = np.tile(_f32([0, 0, 1]), (4, 1))
    stride = 2
    idx = np.array([0, 1, stride + 1, stride + 1, stride, 0], dtype=np.uint32)
    return dict(vertices=verts, normals=normals, indices=idx, transform=translate(position), type="Plane")


def create_sphere(position):
    """Primitives::createSphere (src/geometry/Primitives.cpp:63-140): r = 0.3, tessU 180, tessV 90."""
    tessU, tessV, radius = 180, 90, np.float32(0.3)
    phi_step = np.float32(2.0) * np.float32(np.pi) / np.float32(tessU)
    theta_step = np.float32(np.pi) / np.float32(tessV - 1)
    lat = np.arange(tessV, dtype=np.float32) * theta_step
    lon = np.arange(tessU + 1, dtype=np.float32) * phi_step
    st, ct = np.sin(lat, dtype=np.float32), np.cos(lat, dtype=np.float32)
    sp, cp = np.sin(lon, dtype=np.float32), np.cos(lon, dtype=np.float32)
    normals = np.stack([np.outer(st, cp), np.outer(ct, np.ones_like(cp)), np.outer(st, sp)], axis=-1).astype(np.float32).reshape(-1, 3)
    verts = (normals * radius).astype(np.float32)
    cols = tessU + 1
    la, lo = np.meshgrid(np.arange(tessV - 1), np.arange(tessU), indexing="ij")
    a = la * cols + lo
    idx = np.stack([a, a + 1, a + cols + 1, a + cols + 1, a + cols, a], axis=-1).astype(np.uint32).reshape(-1)
    return dict(vertices=verts, normals=normals, indices=idx, transform=translate(position), type="Sphere")
