"""Shared helpers for the -m gpu parity tests: run the CUDA path through the C ABI and the oracle on the same inputs."""
import numpy as np


class GpuFrame:
    pass


def gpu_render(ctx, grt, rec, P, *, meshes=None, trace_rect=None, trace_max=0, counters=False, tile=None):
    """Upload, build, launch through the C ABI; returns rgb8, rgb32f, traces, counters."""
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    ctx.validate_accel(0)
    P.handle, P.d_particles = handle, dpart
    if meshes is not None:
        ctx.set_meshes(meshes)
        mh, dm = ctx.build_mesh_accel()
        ctx.validate_accel(1)
        P.mesh_handle, P.d_meshes = mh, dm
    else:
        ctx.set_meshes([])
        ctx.build_mesh_accel()
        P.mesh_handle, P.d_meshes = 0, None
    W, H = P.width, P.height
    d_out = ctx.malloc(W * H * 3)
    d_f32 = ctx.malloc(W * H * 3 * 4)
    ctx.memset(d_out, 0, W * H * 3)
    ctx.memset(d_f32, 0, W * H * 3 * 4)
    d_ids = d_cnt = None
    if trace_rect and trace_max:
        npx = trace_rect[2] * trace_rect[3]
        d_ids = ctx.malloc(npx * trace_max * 4)
        d_cnt = ctx.malloc(npx * 4)
        ctx.memset(d_ids, 0xFF, npx * trace_max * 4)
        ctx.memset(d_cnt, 0, npx * 4)
    ctx.set_debug_outputs(d_f32, d_ids, d_cnt, trace_rect or (0, 0, 0, 0), trace_max)
    ctx.enable_counters(counters)
    if counters:
        ctx.counters(reset=True)
    P.output_buffer = d_out
    if tile:
        ctx.launch_tile(P, *tile)
    else:
        ctx.launch(P)
    ctx.sync()
    out = GpuFrame()
    out.rgb8 = ctx.to_host(d_out, (H, W, 3), np.uint8)
    out.rgb32f = ctx.to_host(d_f32, (H, W, 3), np.float32)
    out.ms = ctx.last_launch_ms()
    if d_ids:
        npx = trace_rect[2] * trace_rect[3]
        out.trace_ids = ctx.to_host(d_ids, (npx, trace_max), np.uint32)
        out.trace_counts = ctx.to_host(d_cnt, (npx,), np.uint32)
        ctx.free(d_ids)
        ctx.free(d_cnt)
    out.counters = ctx.counters() if counters else None
    ctx.set_debug_outputs(None, None, None, (0, 0, 0, 0), 0)
    ctx.enable_counters(False)
    ctx.free(d_out)
    ctx.free(d_f32)
    P.output_buffer = None
    return out


def compare(gpu, ora, *, tol=1e-3, max_flag_frac=0.01, label=""):
    """north_star acceptance: float RGB max-abs <= 1e-3 and PSNR >= 50 dB, RGB8 within 1 LSB.
    Pixels the oracle flags tie/threshold-ambiguous are reported and excluded from the max-abs."""
    from conftest import psnr
    flags = ora["flags"].astype(bool)
    diff = np.abs(gpu.rgb32f - ora["rgb32f"]).max(axis=2)
    clean = diff[~flags]
    res = dict(max_abs=float(clean.max()) if clean.size else 0.0, max_abs_all=float(diff.max()), flagged=int(flags.sum()),
               flag_frac=float(flags.mean()), psnr=psnr(np.clip(gpu.rgb32f, 0, 1), np.clip(ora["rgb32f"], 0, 1)),
               lsb=int(np.abs(gpu.rgb8.astype(np.int32) - ora["rgb8"].astype(np.int32))[~flags].max()) if clean.size else 0,
               n_bad=int((clean > tol).sum()))
    print(f"[parity {label}] {res}")
    assert res["flag_frac"] <= max_flag_frac, res
    assert res["max_abs"] <= tol, res
    assert res["psnr"] >= 50.0, res
    assert res["lsb"] <= 1, res
    return res
