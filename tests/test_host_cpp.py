"""The C++ host mirror (gaussian-ray-tracing_b200/host: GaussianData, Camera, Primitives, GaussianTracer) through its
C test entry points.  CPU part: against the Python mirror and, when oracle/_ref is built, against the REFERENCE's own
GaussianData.cpp / Camera.cpp.  GPU part: the facade's whole call sequence renders the same frame as the Python path."""
import ctypes as C
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def host(grt):
    path = grt.LIBHOST_PATH
    assert os.path.exists(path), f"{path} missing: run make -C gaussian-ray-tracing_b200"
    lib = C.CDLL(path)
    lib.grtHostLoadPly.restype = C.c_void_p
    lib.grtHostLoadPly.argtypes = [C.c_char_p, C.POINTER(C.c_ulonglong), C.POINTER(C.c_float)]
    lib.grtHostFree.argtypes = [C.c_void_p]
    return lib


def _load(host, path):
    n = C.c_ulonglong(0)
    center = (C.c_float * 3)()
    p = host.grtHostLoadPly(path.encode(), C.byref(n), center)
    if not p:
        return None, None
    rec = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(n.value, 59)).copy()
    host.grtHostFree(p)
    return rec, np.array(list(center), np.float32)


def test_cpp_ply_loader_matches_python_and_reference(host, grt, orc, tmp_path):
    props = grt.synth_ply_properties(5000, 77)
    path = str(tmp_path / "a.ply")
    grt.write_ply(path, props)
    rec, center = _load(host, path)
    mine = grt.activate(props)
    assert rec.shape == mine.shape
    assert np.abs(rec - mine).max() < 2e-6  # glibc expf vs numpy exp: <= 1 ulp
    exact = [0, 1, 2] + list(range(11, 59))  # positions and SH are copies
    assert np.array_equal(rec[:, exact], mine[:, exact])
    assert np.allclose(center, grt.scene.gaussian_center(mine), atol=1e-6)
    if orc.ref_available():  # the reference's own loader: same libm, so bit-identical records and centre
        ref, rcenter = orc.ref_load_ply(path)
        assert np.array_equal(ref, rec)
        assert np.array_equal(np.asarray(rcenter, np.float32), center)


def test_cpp_ply_loader_formats_and_errors(host, grt, tmp_path):
    props = grt.synth_ply_properties(40, 78)
    names = list(grt.scene.PLY_PROPERTIES)
    # ascii, with an extra leading double column and shuffled property order
    order = np.random.default_rng(0).permutation(len(names))
    path = str(tmp_path / "ascii.ply")
    with open(path, "w") as f:
        f.write("ply\nformat ascii 1.0\nelement vertex 40\nproperty double extra\n")
        f.write("".join(f"property float {names[j]}\n" for j in order) + "end_header\n")
        for row in props:
            f.write("1.5 " + " ".join(repr(float(row[j])) for j in order) + "\n")
    rec, _ = _load(host, path)
    assert np.abs(rec - grt.activate(props)).max() < 2e-6
    # big endian
    path = str(tmp_path / "be.ply")
    with open(path, "wb") as f:
        f.write(("ply\nformat binary_big_endian 1.0\nelement vertex 40\n" + "".join(f"property float {n}\n" for n in names) + "end_header\n").encode())
        f.write(props.astype(">f4").tobytes())
    rec, _ = _load(host, path)
    assert np.abs(rec - grt.activate(props)).max() < 2e-6
    # a missing property is an error (happly throws, happly.h:857), so is a missing file
    path = str(tmp_path / "missing.ply")
    keep = [i for i, n in enumerate(names) if n != "f_rest_44"]
    grt.write_ply(path, props[:, keep], names=[names[i] for i in keep])
    assert _load(host, path)[0] is None
    assert _load(host, str(tmp_path / "nope.ply"))[0] is None
    # empty scene
    path = str(tmp_path / "empty.ply")
    grt.write_ply(path, props[:0])
    rec, _ = _load(host, path)
    assert rec.shape == (0, 59)


def test_cpp_camera_and_primitives(host, grt, orc):
    f3 = C.c_float * 3
    for eye, look, fov, asp in [((0, 0, 3), (0.01, -0.2, 0.03), 60.0, 16 / 9), ((1.5, 0.5, -2), (0, 0, 0), 45.0, 1.0)]:
        U, V, W = f3(), f3(), f3()
        host.grtHostCameraUVW(f3(*eye), f3(*look), f3(0, 1, 0), C.c_float(fov), C.c_float(asp), U, V, W)
        Up, Vp, Wp = grt.camera_uvw(eye, look, (0, 1, 0), fov, np.float32(asp))
        assert np.allclose(list(U), Up, rtol=2e-6, atol=1e-7) and np.allclose(list(V), Vp, rtol=2e-6, atol=1e-7) and np.array_equal(np.float32(list(W)), Wp)
        if orc.ref_available():
            Ur, Vr, Wr = orc.ref_camera_uvw(eye, look, (0, 1, 0), fov, asp)
            assert np.array_equal(np.float32(list(U)), Ur) and np.array_equal(np.float32(list(V)), Vr) and np.array_equal(np.float32(list(W)), Wr)
    for kind, make in [(0, grt.create_plane), (1, grt.create_sphere)]:
        nv, ni = C.c_uint(0), C.c_uint(0)
        pos = f3(0.25, -0.5, 1.75)
        host.grtHostPrimitive(kind, pos, C.byref(nv), C.byref(ni), None, None, None, None)
        v = np.zeros((nv.value, 3), np.float32); n = np.zeros((nv.value, 3), np.float32)
        idx = np.zeros(ni.value, np.uint32); T = np.zeros(16, np.float32)
        host.grtHostPrimitive(kind, pos, C.byref(nv), C.byref(ni), v.ctypes.data_as(C.POINTER(C.c_float)), n.ctypes.data_as(C.POINTER(C.c_float)),
                              idx.ctypes.data_as(C.POINTER(C.c_uint)), T.ctypes.data_as(C.POINTER(C.c_float)))
        ref = make([0.25, -0.5, 1.75])
        assert (nv.value, ni.value) == (len(ref["vertices"]), len(ref["indices"]))
        assert (nv.value, ni.value) == ((4, 6) if kind == 0 else (16290, 96120))
        assert np.array_equal(idx, ref["indices"])
        assert np.abs(v - ref["vertices"]).max() < 1e-6 and np.abs(n - ref["normals"]).max() < 1e-6
        assert np.array_equal(T, ref["transform"].reshape(16))


@pytest.mark.gpu
@pytest.mark.parametrize("flags,mesh_type", [(0, 0), (1, 0), (6, 0), (6, 2)])
def test_cpp_facade_renders_like_python_facade(host, grt, ctx, tmp_path, flags, mesh_type):
    """GaussianTracer(ply) -> setSize -> initializeOptix -> updateCamera -> [createPlane/createSphere] -> render."""
    props = grt.synth_ply_properties(4000, 79)
    path = str(tmp_path / "s.ply")
    grt.write_ply(path, props)
    W, H = 96, 64
    out = np.zeros((H, W, 3), np.uint8)
    ms = C.c_float(0)
    err = C.create_string_buffer(512)
    rc = host.grtHostRenderPly(path.encode(), W, H, 3, flags, mesh_type, out.ctypes.data_as(C.POINTER(C.c_ubyte)), C.byref(ms), err, 512)
    assert rc == 0, err.value.decode()
    tr = grt.GaussianTracer(path, ctx=ctx)
    tr.setSize(W, H)
    tr.initializeOptix()
    cam = grt.Camera()
    cam.setEye([0, 0, 3]); cam.setLookat(tr.getGaussianCenter()); cam.setUp([0, 1, 0]); cam.setFovY(60.0)
    tr.updateCamera(cam, True)
    tr.params.sh_degree_max = 3
    tr.params.mode_fisheye = flags & 1
    tr.setRenderType(mesh_type)
    if flags & 2:
        tr.createPlane()
    if flags & 4:
        tr.createSphere()
    buf = grt.CUDAOutputBuffer(ctx, W, H)
    tr.render(buf)
    img = buf.download()
    d = np.abs(img.astype(int) - out.astype(int))
    # same kernels, same records; only libm-vs-numpy ulp differences in the host-side setup can differ
    assert (d > 1).mean() < 0.002 and out.max() > 0, (d.max(), (d > 0).mean())
    assert ms.value > 0


def _read_png_rgb8(path):
    """Minimal PNG reader for the test: 8-bit RGB, non-interlaced, filter type 0 rows; checks every chunk CRC."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(typ + body) & 0xFFFFFFFF
        if typ == b"IHDR":
            w, h, depth, ctype, comp, flt, inter = struct.unpack(">IIBBBBB", body)
            assert (depth, ctype, comp, flt, inter) == (8, 2, 0, 0, 0)
        elif typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)   # zlib verifies the Adler-32
    assert (raw[:, 0] == 0).all()
    return raw[:, 1:].reshape(h, w, 3)


@pytest.mark.gpu
def test_cpp_frame_sinks_png_and_ppm(host, grt, tmp_path, monkeypatch):
    """Headless frame sink: the PNG and PPM files hold exactly the frame the facade returned."""
    props = grt.synth_ply_properties(3000, 80)
    path = str(tmp_path / "s.ply")
    grt.write_ply(path, props)
    W, H = 300, 250   # 226 KB of scanlines: more than one 64 KB stored block
    for ext in ("png", "ppm"):
        sink = str(tmp_path / f"frame.{ext}")
        monkeypatch.setenv("GRT_HOST_FRAME_SINK", sink)
        out = np.zeros((H, W, 3), np.uint8)
        err = C.create_string_buffer(512)
        rc = host.grtHostRenderPly(path.encode(), W, H, 3, 8, 0, out.ctypes.data_as(C.POINTER(C.c_ubyte)), None, err, 512)
        assert rc == 0, err.value.decode()
        assert out.max() > 0
        if ext == "png":
            img = _read_png_rgb8(sink)
        else:
            blob = open(sink, "rb").read()
            hdr = f"P6\n{W} {H}\n255\n".encode()
            assert blob.startswith(hdr)
            img = np.frombuffer(blob[len(hdr):], np.uint8).reshape(H, W, 3)
        assert np.array_equal(img, out)
