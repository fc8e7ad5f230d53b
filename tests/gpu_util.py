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


FLAG_TIE, FLAG_ALPHA, FLAG_TRANSMIT, FLAG_MESH = 1, 2, 4, 8


def parity_stats(gpu_f32, gpu_u8, ora_f32, ora_u8, flags, *, tol=1e-3, label="", max_tie_frac=0.01, max_threshold_frac=2e-3):
    """north_star acceptance on a frame (or band): float RGB max-abs <= tol and PSNR >= 50 dB, RGB8 within 1 LSB, asserted on
    EVERY pixel except those the oracle flags as equal-t ties (FLAG_TIE; FLAG_MESH = the same for two triangles) -- the one
    freedom north_star grants.  A pixel flagged only because some alpha came within 5e-7 of alpha_min or some T within 5e-8 of
    minTransmittance (a discontinuous decision that a last-ulp difference of expf flips) has to pass too, unless the decision
    really flipped: such pixels are counted and reported separately and may not exceed max_threshold_frac of the frame."""
    from conftest import psnr
    flags = np.asarray(flags)
    tie = (flags & (FLAG_TIE | FLAG_MESH)) != 0
    thr = ((flags & (FLAG_ALPHA | FLAG_TRANSMIT)) != 0) & ~tie
    diff = np.abs(gpu_f32 - ora_f32).max(axis=2)
    lsb = np.abs(gpu_u8.astype(np.int32) - ora_u8.astype(np.int32)).max(axis=2)
    bad = (diff > tol) | (lsb > 1)
    res = dict(max_abs_clean=float(diff[~tie & ~thr].max()) if (~tie & ~thr).any() else 0.0,
               max_abs_nontie=float(diff[~tie].max()) if (~tie).any() else 0.0, max_abs_all=float(diff.max()),
               tie_frac=float(tie.mean()), threshold_frac=float(thr.mean()), threshold_flipped=int((bad & thr).sum()),
               flag_mix={"tie": int(((flags & FLAG_TIE) != 0).sum()), "alpha": int(((flags & FLAG_ALPHA) != 0).sum()),
                         "transmit": int(((flags & FLAG_TRANSMIT) != 0).sum()), "mesh": int(((flags & FLAG_MESH) != 0).sum())},
               psnr=psnr(np.clip(gpu_f32, 0, 1), np.clip(ora_f32, 0, 1)), lsb_nontie=int(lsb[~tie & ~(bad & thr)].max()) if (~tie).any() else 0,
               n_bad=int((bad & ~tie & ~thr).sum()), pixels=int(flags.size))
    print(f"[parity {label}] {res}")
    assert res["n_bad"] == 0, res                                  # every unflagged pixel is inside the tolerance
    assert res["threshold_flipped"] <= max(1, int(max_threshold_frac * flags.size)), res
    assert res["tie_frac"] <= max_tie_frac, res
    assert res["psnr"] >= 50.0, res
    assert res["lsb_nontie"] <= 1, res
    return res


def compare(gpu, ora, *, tol=1e-3, max_flag_frac=0.01, label=""):
    """Whole-frame form of parity_stats for a GpuFrame and an oracle render."""
    return parity_stats(gpu.rgb32f, gpu.rgb8, ora["rgb32f"], ora["rgb8"], ora["flags"], tol=tol, label=label, max_tie_frac=max_flag_frac)


def check_traces(gpu, ora, rect, flags, trace_max, *, min_compared=1):
    """Consumed particle-ID sequences pixel by pixel: identical except where the oracle flags an equal-t tie."""
    x0, y0, w, h = rect
    n_cmp = n_bad = n_multi = 0
    for j in range(h):
        for i in range(w):
            slot = j * w + i
            if flags[y0 + j, x0 + i] & (FLAG_TIE | FLAG_MESH):
                continue
            n_cmp += 1
            c = int(ora["trace_counts"][slot])
            assert int(gpu.trace_counts[slot]) == c, (i, j, int(gpu.trace_counts[slot]), c)
            k = min(c, trace_max)
            if not np.array_equal(gpu.trace_ids[slot, :k], ora["trace_ids"][slot, :k]):
                n_bad += 1
    print(f"[traces] compared {n_cmp} pixels, {n_bad} sequence mismatches")
    assert n_cmp >= min_compared and n_bad == 0
    return n_cmp
