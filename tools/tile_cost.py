#!/usr/bin/env python
"""Cost map of the bench frame: times grtLaunchTile over a grid of pixel blocks (finds straggler regions)."""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
grt = importlib.import_module("gaussian-ray-tracing_b200")
W, H = 1920, 1080
rec = grt.activate(grt.synth_ply_properties(1_000_000, 1002))
ctx = grt.Context(0)
ctx.set_particles(rec, 0.01)
handle, dpart = ctx.build_particle_accel()
cam = grt.default_camera(rec)
U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3)
P.handle, P.d_particles = handle, dpart
d_out = ctx.malloc(W * H * 3)
P.output_buffer = d_out
ctx.launch(P); ctx.sync()
print("full frame ms", ctx.last_launch_ms())
bw, bh = int(sys.argv[1]) if len(sys.argv) > 1 else 128, int(sys.argv[2]) if len(sys.argv) > 2 else 72
rows = []
for y0 in range(0, H, bh):
    row = []
    for x0 in range(0, W, bw):
        w, h = min(bw, W - x0), min(bh, H - y0)
        ctx.launch_tile(P, x0, y0, w, h); ctx.sync()
        row.append(ctx.last_launch_ms())
    rows.append(row)
a = np.array(rows)
print("sum of block ms", a.sum(), "max", a.max(), "at", np.unravel_index(a.argmax(), a.shape))
np.set_printoptions(linewidth=250, precision=2, suppress=True)
print(a)
ctx.close()
