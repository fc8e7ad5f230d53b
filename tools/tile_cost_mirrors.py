#!/usr/bin/env python
"""Cost map (ms per 120x120 block, each launched alone) of BASELINE configs[3] (mirror plane + sphere) next to the plain scene."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
grt = importlib.import_module("gaussian-ray-tracing_b200")
W, H = 1920, 1080
rec = grt.activate(grt.synth_ply_properties(1_000_000, 1002))
ctx = grt.Context(0)
ctx.set_wavefront(False)
ctx.set_particles(rec, 0.01)
handle, dpart = ctx.build_particle_accel()
cam = grt.default_camera(rec)
U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3)
P.handle, P.d_particles = handle, dpart
d_out = ctx.malloc(W * H * 3)
P.output_buffer = d_out
np.set_printoptions(linewidth=250, precision=2, suppress=True)


def cost_map(tag):
    rows = []
    for y0 in range(0, H, 120):
        row = []
        for x0 in range(0, W, 120):
            ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
            ctx.enable_counters(True); ctx.counters(reset=True)
            ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
            c = ctx.counters(); ctx.enable_counters(False)
            ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
            row.append((ctx.last_launch_ms(), c["segments"] / c["rays"], c["warp_nodes"] / (c["rays"] / 32), c["warp_leaf_prims"] / (c["rays"] / 32)))
        rows.append(row)
    a = np.array(rows)
    print(tag, "ms"); print(a[:, :, 0])
    print(tag, "segments/ray"); print(a[:, :, 1])
    print(tag, "nodes/packet"); print(a[:, :, 2].round(0))
    print(tag, "records/packet"); print(a[:, :, 3].round(0))


plane = grt.create_plane([0, 0, 0])
plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
ctx.set_meshes([plane, grt.create_sphere([0.45, 0.0, 1.8])])
P.mesh_handle, P.d_meshes = ctx.build_mesh_accel()
P.type = 0
cost_map("mirrors")
ctx.close()
