#!/usr/bin/env python
"""Profiling driver: the bench workload (1M Gaussians, 1080p, deg 3) with a counted pass, then a few plain launches.
Used under ncu (--set full -k regex:trace_kernel) and on its own for the traversal counters.
--mirrors adds configs[3]'s mirror plane + sphere; --cache keeps the activated scene in /tmp between runs (tuning builds
of tools/run_variants.sh share it); the frame hashes (RGB8 and float RGB) tell whether two builds render the same bits."""
import argparse
import hashlib
import importlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--seed", type=int, default=1002)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--deg", type=int, default=3)
    ap.add_argument("--fisheye", action="store_true")
    ap.add_argument("--mirrors", action="store_true")
    ap.add_argument("--mesh-type", type=int, default=0)
    ap.add_argument("--launches", type=int, default=3)
    ap.add_argument("--no-counters", action="store_true")
    ap.add_argument("--cache", action="store_true")
    ap.add_argument("--no-wavefront", action="store_true")
    a = ap.parse_args()
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    cache = f"/tmp/grt_scene_{a.n}_{a.seed}.npy"
    if a.cache and os.path.exists(cache):
        rec = np.load(cache)
    else:
        rec = grt.activate(grt.synth_ply_properties(a.n, a.seed))
        if a.cache:
            np.save(cache, rec)
    ctx = grt.Context(0)
    ctx.set_wavefront(not a.no_wavefront)
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    cam = grt.default_camera(rec)
    U, V, W = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(a.width) / np.float32(a.height)))
    P = grt.make_params(a.width, a.height, cam["eye"], U, V, W, sh_degree_max=a.deg, fisheye=a.fisheye)
    P.handle, P.d_particles = handle, dpart
    if a.mirrors:
        plane = grt.create_plane([0, 0, 0])
        plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
        sphere = grt.create_sphere([0.45, 0.0, 1.8])
        ctx.set_meshes([plane, sphere])
        mh, dm = ctx.build_mesh_accel()
        P.mesh_handle, P.d_meshes, P.type = mh, dm, a.mesh_type
    npix = a.width * a.height
    d_out = ctx.malloc(npix * 3)
    d_f32 = ctx.malloc(npix * 12)
    P.output_buffer = d_out
    out = {"build": ctx.build_info()}
    if not a.no_counters:
        ctx.enable_counters(True)
        ctx.counters(reset=True)
        ctx.launch(P)
        ctx.sync()
        c = ctx.counters()
        ctx.enable_counters(False)
        rays = max(c["rays"], 1)
        out["counters"] = c
        out["per_ray"] = {k: c[k] / rays for k in ("segments", "rounds", "hits_gathered", "hits_shaded")}
        warps = rays / 32
        out["per_warp"] = {k: c[k] / warps for k in ("warp_nodes", "warp_leaf_prims", "warp_rounds", "warp_prims_hit", "warp_consume_iters",
                                                       "warp_candidate_iters", "warp_composite_rounds", "warp_loop_iters", "warp_forced_flushes", "warp_list_full")}
        d = c.get("dbg", [0] * 8)
        out["per_ray"].update({"inserts": d[0] / rays, "shift_steps": d[1] / rays, "ring_cnt_at_insert": d[2] / max(d[0], 1),
                               "candidates": d[3] / rays, "proxy_hits": d[4] / rays, "spills": d[5] / rays, "second_face_same_round": d[6] / rays, "second_face_last8": d[7] / rays})
    ms = []
    for _ in range(a.launches):
        ctx.launch(P)
        ctx.sync()
        ms.append(ctx.last_launch_ms())
    out["launch_ms"] = ms
    out["mrays_s"] = npix / (min(ms) * 1e-3) / 1e6
    # one more launch with the float frame attached: hashes of both outputs
    ctx.set_debug_outputs(d_rgb32f=d_f32)
    ctx.launch(P)
    ctx.sync()
    ctx.set_debug_outputs()
    out["rgb8_sha1"] = hashlib.sha1(ctx.to_host(d_out, (npix * 3,), np.uint8).tobytes()).hexdigest()[:16]
    out["rgb32f_sha1"] = hashlib.sha1(ctx.to_host(d_f32, (npix * 3,), np.float32).tobytes()).hexdigest()[:16]
    print(json.dumps(out, default=float))
    ctx.free(d_out)
    ctx.free(d_f32)
    ctx.close()


if __name__ == "__main__":
    main()
