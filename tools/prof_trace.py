#!/usr/bin/env python
"""Profiling driver: the bench workload (1M Gaussians, 1080p, deg 3) with a counted pass, then a few plain launches.
Used under ncu (--set full -k regex:trace_kernel) and on its own for the traversal counters."""
import argparse
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
    ap.add_argument("--launches", type=int, default=3)
    ap.add_argument("--no-counters", action="store_true")
    a = ap.parse_args()
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    rec = grt.activate(grt.synth_ply_properties(a.n, a.seed))
    ctx = grt.Context(0)
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    cam = grt.default_camera(rec)
    U, V, W = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(a.width) / np.float32(a.height)))
    P = grt.make_params(a.width, a.height, cam["eye"], U, V, W, sh_degree_max=a.deg, fisheye=a.fisheye)
    P.handle, P.d_particles = handle, dpart
    d_out = ctx.malloc(a.width * a.height * 3)
    P.output_buffer = d_out
    out = {"build": ctx.build_info()}
    if not a.no_counters:
        ctx.enable_counters(True)
        ctx.counters(reset=True)
        ctx.launch(P)
        ctx.sync()
        c = ctx.counters()
        ctx.enable_counters(False)
        rays = c["rays"]
        out["counters"] = c
        out["per_ray"] = {k: c[k] / rays for k in ("segments", "rounds", "hits_gathered", "hits_shaded")}
        warps = rays / 32
        out["per_warp"] = {k: c[k] / warps for k in ("warp_nodes", "warp_leaf_prims", "warp_rounds")}
    ms = []
    for _ in range(a.launches):
        ctx.launch(P)
        ctx.sync()
        ms.append(ctx.last_launch_ms())
    out["launch_ms"] = ms
    out["mrays_s"] = a.width * a.height / (min(ms) * 1e-3) / 1e6
    print(json.dumps(out, default=float))
    ctx.free(d_out)
    ctx.close()


if __name__ == "__main__":
    main()
