#!/usr/bin/env python
"""Kernel time of every BASELINE.json config on one GPU (configs[1..4]; configs[0] is the oracle's parity case).
CUDA-event time per launch (grtLastLaunchMs), 3 warm-up + median of N frames, 256 MB L2-flush write between frames.
One JSON object per config on stdout; the headline number stays bench.py's."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
grt = importlib.import_module("gaussian-ray-tracing_b200")


def frames(ctx, P, n, flush):
    ms = []
    for i in range(3 + n):
        ctx.memset(flush, i & 0xFF, 256 << 20)
        ctx.launch(P)
        ctx.sync()
        if i >= 3:
            ms.append(ctx.last_launch_ms())
    return float(np.median(ms)), float(min(ms)), float(max(ms))


def counted(ctx, P):
    ctx.enable_counters(True); ctx.counters(reset=True)
    ctx.launch(P); ctx.sync()
    c = ctx.counters(); ctx.enable_counters(False)
    return c


def main():
    nframes = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    ctx = grt.Context(0)
    flush = ctx.malloc(256 << 20)
    W, H = 1920, 1080
    rec = grt.activate(grt.synth_ply_properties(1_000_000, 1002))
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    cam = grt.default_camera(rec)
    U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
    d_out = ctx.malloc(W * H * 3)
    out = []
    for name, fisheye, mirrors in [("configs[1] 1M 1080p pinhole", False, False), ("configs[2] 1M 1080p fisheye", True, False),
                                   ("configs[3] 1M 1080p pinhole + mirror plane and sphere", False, True)]:
        P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3, fisheye=fisheye)
        P.handle, P.d_particles, P.output_buffer = handle, dpart, d_out
        if mirrors:
            plane = grt.create_plane([0, 0, 0])
            plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
            ctx.set_meshes([plane, grt.create_sphere([0.45, 0.0, 1.8])])
            P.mesh_handle, P.d_meshes = ctx.build_mesh_accel()
            P.type = 0
        c = counted(ctx, P)
        med, lo, hi = frames(ctx, P, nframes, flush)
        out.append({"config": name, "ms_per_frame": med, "ms_min": lo, "ms_max": hi, "Mrays_s": W * H / med / 1e3,
                    "Msegments_s": c["segments"] / med / 1e3, "segments_per_ray": c["segments"] / c["rays"],
                    "hits_gathered_per_ray": c["hits_gathered"] / c["rays"], "hits_shaded_per_ray": c["hits_shaded"] / c["rays"],
                    "bvh_build_ms": ctx.build_info()["build_ms"], "mesh_build_ms": ctx.build_info()["mesh_build_ms"]})
        print(json.dumps(out[-1]), flush=True)
        if mirrors:
            ctx.set_meshes([]); ctx.build_mesh_accel()
    ctx.free(d_out)
    if os.environ.get("GRT_SKIP_CONFIG4"):
        return
    # configs[4]: 6M Gaussians, 3840x2160, 64-camera ring (all 64 on this one GPU; bench.py --gpus N partitions them)
    W, H = 3840, 2160
    rec = grt.activate(grt.synth_ply_properties(6_000_000, 1003))
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    cam = grt.default_camera(rec)
    d_out = ctx.malloc(W * H * 3)
    ms, hg = [], []
    for k in range(64):
        th = 2 * np.pi * k / 64
        eye = np.array([3 * np.sin(th), 0.5, 3 * np.cos(th)], np.float32)
        U, V, Wv = grt.camera_uvw(eye, cam["lookat"], cam["up"], 60.0, float(np.float32(W) / np.float32(H)))
        P = grt.make_params(W, H, eye, U, V, Wv, sh_degree_max=3)
        P.handle, P.d_particles, P.output_buffer = handle, dpart, d_out
        if k == 0:
            ctx.launch(P); ctx.sync()  # warm-up
            c = counted(ctx, P)
            hg = c["hits_gathered"] / c["rays"]
        ctx.memset(flush, k, 256 << 20)
        ctx.launch(P); ctx.sync()
        ms.append(ctx.last_launch_ms())
    tot = float(sum(ms))
    out.append({"config": "configs[4] 6M 4K, 64-camera ring on one GPU", "ms_all_64_cameras": tot, "ms_per_frame": tot / 64,
                "ms_min": float(min(ms)), "ms_max": float(max(ms)), "Mrays_s": 64 * W * H / tot / 1e3, "hits_gathered_per_ray_cam0": hg,
                "bvh_build_ms": ctx.build_info()["build_ms"], "n_live": ctx.build_info()["n_live"]})
    print(json.dumps(out[-1]), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
