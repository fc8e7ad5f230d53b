"""-m gpu: the CUDA hot path (through the C ABI) against the CPU oracle on identical inputs.

Acceptance (BASELINE.json north_star): per-pixel float RGB within 1e-3 max-abs and >= 50 dB PSNR, RGB8 within
1 LSB, consumed particle-ID sequences identical except at equal-t ties (pixels the oracle flags)."""
import numpy as np
import pytest

from conftest import make_scene
from gpu_util import check_traces, compare, gpu_render

pytestmark = pytest.mark.gpu


def _oracle(orc, rec, P, meshes=None, **kw):
    S = orc.Scene(rec, 0.01)
    if meshes:
        S.set_meshes(meshes)
    return S.render(orc.params_from_grt(P), K=7, **kw)


def _check_traces(gpu, ora, rect, W, flags, trace_max):
    return check_traces(gpu, ora, rect, flags, trace_max)


def _trace_pixels(rect, width):
    return np.array([(rect[1] + j) * width + rect[0] + i for j in range(rect[3]) for i in range(rect[2])], np.uint32)


@pytest.mark.parametrize("deg", [0, 1, 2, 3])
def test_pinhole_parity_small(ctx, grt, orc, deg):
    rec, P, _ = make_scene(grt, 20_000, 101, 160, 120, deg=deg)
    rect = (60, 40, 32, 24)
    tp = np.array([(rect[1] + j) * P.width + rect[0] + i for j in range(rect[3]) for i in range(rect[2])], np.uint32)
    ora = _oracle(orc, rec, P, trace_pixels=tp, trace_max=96)
    gpu = gpu_render(ctx, grt, rec, P, trace_rect=rect, trace_max=96, counters=True)
    compare(gpu, ora, label=f"pinhole deg{deg}")
    _check_traces(gpu, ora, rect, P.width, ora["flags"], 96)
    c, s = gpu.counters, ora["stats"]
    assert c["stack_overflow"] == 0
    assert c["rays"] == s["rays"]
    # h_g / h_s are K-independent up to ties (the GPU streams hits through 32-entry rings, the oracle gathers 7 per round)
    assert abs(c["hits_gathered"] - s["hits_gathered"]) <= 4 * max(1, s["flagged"]), (c, s)
    assert abs(c["hits_shaded"] - s["hits_shaded"]) <= 4 * max(1, s["flagged"]), (c, s)


def test_config1_100k_800x800(ctx, grt, orc):
    """BASELINE.json configs[0]: 100k synthetic Gaussians, SH deg 3, 800x800 pinhole, 1 frame."""
    rec, P, _ = make_scene(grt, 100_000, 1001, 800, 800, deg=3)
    ora = _oracle(orc, rec, P)
    gpu = gpu_render(ctx, grt, rec, P)
    res = compare(gpu, ora, label="config1")
    print("config1 gpu ms", gpu.ms, "oracle s", ora["stats"]["seconds"], "threads", ora["stats"]["threads"], res)


def test_fisheye_parity(ctx, grt, orc):
    rec, P, _ = make_scene(grt, 20_000, 103, 160, 160, deg=3, fisheye=True)
    # hit-ID sequences on a block at the centre (directions change sign inside packets: the incoherent node test) and one
    # at the rim of the fisheye circle (strongly divergent rays, pixels with r > 1 next to traced ones)
    for rect in [(64, 68, 32, 24), (0, 60, 40, 24)]:
        ora = _oracle(orc, rec, P, trace_pixels=_trace_pixels(rect, P.width), trace_max=96)
        gpu = gpu_render(ctx, grt, rec, P, trace_rect=rect, trace_max=96, counters=True)
        compare(gpu, ora, label=f"fisheye {rect}", max_flag_frac=0.02)
        check_traces(gpu, ora, rect, ora["flags"], 96, min_compared=200)
        assert gpu.counters["stack_overflow"] == 0 and gpu.counters["order_violations"] == 0
    # r > 1 corners are DEFINED black
    assert gpu.rgb8[0, 0].tolist() == [0, 0, 0] and gpu.rgb8[-1, -1].tolist() == [0, 0, 0]


def _config4_meshes(grt):
    plane = grt.create_plane([0, 0, 0])
    plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
    sphere = grt.create_sphere([0.45, 0.0, 1.8])
    return [plane, sphere]


@pytest.mark.parametrize("mesh_type", [0, 1, 2])
def test_mesh_modes_parity(ctx, grt, orc, mesh_type):
    rec, P, _ = make_scene(grt, 20_000, 104, 192, 128, deg=3, mesh_type=mesh_type)
    meshes = _config4_meshes(grt)
    # hit-ID sequences (all segments of a pixel, in order) on a full-width band through the plane and the sphere: mirror-sphere
    # rim packets fan out and take the per-ray walk (trace_rays_dfs), which must yield the same sequences as the packet mode
    rect, tmax = (0, 40, 192, 48), 160
    ora = _oracle(orc, rec, P, meshes=meshes, trace_pixels=_trace_pixels(rect, P.width), trace_max=tmax)
    gpu = gpu_render(ctx, grt, rec, P, meshes=meshes, trace_rect=rect, trace_max=tmax, counters=True)
    # secondary rays are bit-reproduced too (un-fused ray-path arithmetic), so the same bar applies
    compare(gpu, ora, label=f"mesh type {mesh_type}", max_flag_frac=0.03)
    check_traces(gpu, ora, rect, ora["flags"], tmax, min_compared=5000)
    c = gpu.counters
    assert c["stack_overflow"] == 0 and c["order_violations"] == 0
    print("mesh type", mesh_type, "segments/ray", c["segments"] / c["rays"], "per-ray passes", c["warp_per_ray_passes"])
    if mesh_type != 1:
        assert c["segments"] > c["rays"] and c["warp_per_ray_passes"] > 0  # the divergent (per-ray) mode was exercised
    no_mesh = _oracle(orc, rec, P)
    assert np.abs(no_mesh["rgb32f"] - ora["rgb32f"]).max() > 0.01  # the primitives are actually in view


@pytest.mark.parametrize("mesh_type", [0, 1, 2])
def test_wavefront_equals_single_launch(ctx, grt, mesh_type):
    """Scenes with primitives are traced level by level (wavefront, the default) or by one launch running the whole bounce loop:
    the same packets take the same steps either way, so float frame, RGB8 frame and hit-ID sequences are bit-identical."""
    rec, P, _ = make_scene(grt, 20_000, 104, 192, 128, deg=3, mesh_type=mesh_type)
    meshes = _config4_meshes(grt)
    rect, tmax = (0, 40, 192, 48), 160
    try:
        ctx.set_wavefront(True)
        wf = gpu_render(ctx, grt, rec, P, meshes=meshes, trace_rect=rect, trace_max=tmax, counters=True)
        ctx.set_wavefront(False)
        mono = gpu_render(ctx, grt, rec, P, meshes=meshes, trace_rect=rect, trace_max=tmax, counters=True)
    finally:
        ctx.set_wavefront(True)
    assert np.array_equal(wf.rgb32f.view(np.uint32), mono.rgb32f.view(np.uint32))
    assert np.array_equal(wf.rgb8, mono.rgb8)
    assert np.array_equal(wf.trace_counts, mono.trace_counts) and np.array_equal(wf.trace_ids, mono.trace_ids)
    for k in ("rays", "segments", "hits_gathered", "hits_shaded", "warp_per_ray_passes"):
        assert wf.counters[k] == mono.counters[k], k
    assert wf.counters["stack_overflow"] == 0 and wf.counters["order_violations"] == 0
    if mesh_type != 1:
        assert wf.counters["segments"] > wf.counters["rays"]


def test_mesh_scene_rows_and_tile_equal_full_frame(ctx, grt):
    """The sorted two-launch mode numbers its tiles over the launch's own rows / rectangle: row bands (grtLaunchRows) and an aligned
    tile (grtLaunchTile) of a scene with mirrors reproduce the full frame bit for bit."""
    W, H = 192, 128
    rec, P, _ = make_scene(grt, 20_000, 104, W, H, deg=3, mesh_type=0)
    meshes = _config4_meshes(grt)
    full = gpu_render(ctx, grt, rec, P, meshes=meshes)
    part = gpu_render(ctx, grt, rec, P, meshes=meshes, tile=(32, 40, 96, 48))
    assert np.array_equal(part.rgb8[40:88, 32:128], full.rgb8[40:88, 32:128])
    assert part.rgb8[:40].max() == 0 and part.rgb8[88:].max() == 0
    ctx.set_particles(rec, 0.01)
    P.handle, P.d_particles = ctx.build_particle_accel()
    ctx.set_meshes(meshes)
    P.mesh_handle, P.d_meshes = ctx.build_mesh_accel()
    d = ctx.malloc(W * H * 3)
    P.output_buffer = d
    try:
        for band_rows, G in [(8, 2), (16, 3)]:
            ctx.memset(d, 0, W * H * 3)
            for r in range(G):
                ctx.launch_rows(P, band_rows, G, r)
            ctx.sync()
            assert np.array_equal(ctx.to_host(d, (H, W, 3), np.uint8), full.rgb8), (band_rows, G)
    finally:
        ctx.free(d)
        P.output_buffer = None
        ctx.set_meshes([])
        ctx.build_mesh_accel()


def test_tile_and_batch_equal_full_frame(ctx, grt, orc):
    rec, P, _ = make_scene(grt, 5_000, 105, 100, 70, deg=2)
    full = gpu_render(ctx, grt, rec, P)
    part = gpu_render(ctx, grt, rec, P, tile=(16, 16, 50, 33))
    assert np.array_equal(part.rgb8[16:49, 16:66], full.rgb8[16:49, 16:66])
    assert part.rgb8[:16].max() == 0 and part.rgb8[49:].max() == 0  # untouched outside the tile


@pytest.mark.parametrize("size", [(100, 70), (160, 121)])
def test_row_bands_union_is_frame(ctx, grt, size):
    """grtLaunchRows: G launches with band_stride = G, band_offset = r write disjoint row bands whose union is grtLaunch's frame,
    bit for bit (bands are multiples of the 8-row packet grid, so every packet is the same packet)."""
    W, H = size
    rec, P, _ = make_scene(grt, 5_000, 105, W, H, deg=2)
    full = gpu_render(ctx, grt, rec, P)
    ctx.set_particles(rec, 0.01)
    P.handle, P.d_particles = ctx.build_particle_accel()
    d = ctx.malloc(W * H * 3)
    P.output_buffer = d
    for band_rows, G in [(8, 2), (8, 3), (16, 4), (8, 8), (24, 5)]:
        ctx.memset(d, 0, W * H * 3)
        seen = np.zeros((H, W), bool)
        for r in range(G):
            before = ctx.to_host(d, (H, W, 3), np.uint8)
            ctx.launch_rows(P, band_rows, G, r)
            ctx.sync()
            after = ctx.to_host(d, (H, W, 3), np.uint8)
            own = np.zeros(H, bool)
            for b in range(r, (H + band_rows - 1) // band_rows, G):
                own[b * band_rows:(b + 1) * band_rows] = True
            assert np.array_equal(after[~own], before[~own])        # nothing outside the owned bands is written
            assert not seen[own].any()
            seen[own] = True
        assert seen.all()
        assert np.array_equal(ctx.to_host(d, (H, W, 3), np.uint8), full.rgb8), (band_rows, G)
    with pytest.raises(grt.GrtError):
        ctx.launch_rows(P, 12, 2, 0)      # not a multiple of the packet grid
    with pytest.raises(grt.GrtError):
        ctx.launch_rows(P, 8, 2, 2)       # offset outside the stride
    ctx.free(d)
    P.output_buffer = None


def test_render_host_e2e(ctx, grt, orc):
    rec, P, _ = make_scene(grt, 5_000, 106, 96, 64, deg=3)
    full = gpu_render(ctx, grt, rec, P)
    ctx.set_particles(rec, 0.01)
    P.handle, P.d_particles = ctx.build_particle_accel()
    out = ctx.pinned_empty((64, 96, 3))
    out[:] = 7
    ctx.render_host(P, out)
    assert np.array_equal(out, full.rgb8)


def test_empty_and_bad_handles(ctx, grt):
    rec, P, _ = make_scene(grt, 100, 107, 32, 32)
    ctx.set_particles(rec[:0], 0.01)
    h, _ = ctx.build_particle_accel()
    assert h == 0
    g = gpu_render(ctx, grt, rec[:0], P)
    assert g.rgb8.max() == 0  # nothing to hit: black
    P.handle = 0xDEAD
    d = ctx.malloc(32 * 32 * 3)
    P.output_buffer = d
    with pytest.raises(grt.GrtError):
        ctx.launch(P)
    ctx.free(d)


def test_facade_call_order(ctx, grt, orc, tmp_path):
    """ctor -> setSize -> initializeOptix -> updateCamera -> render, as src/main.cpp:62-100 does."""
    props = grt.synth_ply_properties(3000, 108)
    path = str(tmp_path / "s.ply")
    grt.write_ply(path, props)
    tr = grt.GaussianTracer(path, ctx=ctx)
    tr.setSize(80, 60)
    tr.initializeOptix()
    cam = grt.Camera()
    cam.setEye([0, 0, 3]); cam.setLookat(tr.getGaussianCenter()); cam.setUp([0, 1, 0]); cam.setFovY(60.0)
    tr.updateCamera(cam, True)
    buf = grt.CUDAOutputBuffer(ctx, 80, 60)
    tr.render(buf)                       # default sh_degree_max = 0, like the reference host
    img0 = buf.download()
    tr.params.sh_degree_max = 3          # callers mutate params directly (gui.cpp:190)
    tr.createPlane(); tr.createSphere()
    tr.render(buf)
    img1 = buf.download()
    S = orc.Scene(tr.particles, 0.01)
    S.set_meshes(tr.getPrimitives())
    ora = S.render(orc.params_from_grt(tr.params), K=7)
    ok = ~ora["flags"].astype(bool)
    assert np.abs(img1.astype(int) - ora["rgb8"].astype(int))[ok].max() <= 1
    assert not np.array_equal(img0, img1)
    tr.removePrimitive()
    assert tr.params.mesh_handle == 0


@pytest.mark.parametrize("kind", ["needles_and_disks", "nested_huge_and_tiny"])
def test_adversarial_scenes_parity(ctx, grt, orc, kind):
    """Scenes built to stress the record slots' tight keys and the pending rings: extreme anisotropy with random
    orientation (box keys far too early), and huge transparent shells around dense tiny particles (exit events that
    stay pending across the whole traversal).  Same acceptance bar, plus exact hit-ID sequences on a pixel block."""
    rng = np.random.default_rng(7 if kind == "needles_and_disks" else 8)
    n = 6000
    rec = np.zeros((n, 59), np.float32)
    rec[:, 0:3] = rng.uniform(-0.8, 0.8, (n, 3)) * np.array([1.0, 0.7, 0.8])
    q = rng.normal(size=(n, 4))
    rec[:, 6:10] = q / np.linalg.norm(q, axis=1, keepdims=True)
    if kind == "needles_and_disks":
        s = np.exp(rng.normal(-3.6, 0.5, (n, 3)))
        axis = rng.integers(0, 3, n)
        needle = rng.random(n) < 0.5
        for k in range(3):
            s[:, k] *= np.where(needle, np.where(axis == k, 20.0, 1.0), np.where(axis == k, 0.05, 4.0))
        rec[:, 3:6] = s
        rec[:, 10] = rng.uniform(0.02, 0.9, n)
    else:
        big = rng.random(n) < 0.02
        rec[:, 3:6] = np.where(big[:, None], np.exp(rng.normal(-0.6, 0.2, (n, 3))), np.exp(rng.normal(-4.5, 0.4, (n, 3))))
        rec[:, 10] = np.where(big, rng.uniform(0.011, 0.05, n), rng.uniform(0.3, 0.99, n))
    rec[:, 11:14] = rng.normal(0, 1.0, (n, 3))
    rec[:, 14:59] = rng.normal(0, 0.1, (n, 45))
    W, H = 160, 120
    cam = grt.default_camera(rec)
    U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
    P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3)
    rect, tmax = (72, 52, 16, 12), 160
    tp = np.array([(rect[1] + j) * W + rect[0] + i for j in range(rect[3]) for i in range(rect[2])], np.uint32)
    ora = _oracle(orc, rec, P, trace_pixels=tp, trace_max=tmax)
    gpu = gpu_render(ctx, grt, rec, P, trace_rect=rect, trace_max=tmax, counters=True)
    assert gpu.counters["order_violations"] == 0 and gpu.counters["stack_overflow"] == 0
    compare(gpu, ora, label=kind, max_flag_frac=0.05)
    _check_traces(gpu, ora, rect, W, ora["flags"], tmax)
    print(kind, "hits gathered/ray", gpu.counters["hits_gathered"] / gpu.counters["rays"], "repass", gpu.counters["lanes_repass"],
          "dfs fallbacks", gpu.counters["warp_dfs_fallbacks"])
