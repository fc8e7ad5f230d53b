#!/usr/bin/env python
"""Counters and time of single 120x120 blocks of configs[3] (mirror plane + sphere), each launched alone: what a divergent packet costs.
usage: block_probe.py x0,y0 [x0,y0 ...]"""
import importlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
grt = importlib.import_module("gaussian-ray-tracing_b200")
W, H = 1920, 1080
cache = "/tmp/grt_scene_1000000_1002.npy"
rec = np.load(cache) if os.path.exists(cache) else grt.activate(grt.synth_ply_properties(1_000_000, 1002))
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
plane = grt.create_plane([0, 0, 0])
plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
ctx.set_meshes([plane, grt.create_sphere([0.45, 0.0, 1.8])])
P.mesh_handle, P.d_meshes = ctx.build_mesh_accel()
P.type = 0
for a in sys.argv[1:]:
    x0, y0 = (int(v) for v in a.split(","))
    ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
    ctx.enable_counters(True); ctx.counters(reset=True)
    ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
    c = ctx.counters(); ctx.enable_counters(False)
    ctx.launch_tile(P, x0, y0, 120, 120); ctx.sync()
    ms = ctx.last_launch_ms()
    pr = max(c["warp_per_ray_passes"], 1)
    d = c["dbg"]
    print(json.dumps({"block": [x0, y0], "ms": ms, "segments_per_ray": c["segments"] / c["rays"], "rounds_per_ray": c["rounds"] / c["rays"],
                      "per_ray_passes": c["warp_per_ray_passes"], "dfs_iters_per_pass": c["warp_dfs_iters"] / pr, "repass_lanes": c["lanes_repass"],
                      "inserts_per_ray": d[0] / c["rays"], "shifts_per_insert": d[1] / max(d[0], 1), "proxy_hits_per_ray": d[4] / c["rays"],
                      "spills_per_ray": d[5] / c["rays"], "hits_gathered_per_ray": c["hits_gathered"] / c["rays"], "dfs_fallbacks": c["warp_dfs_fallbacks"],
                      "nodes_per_warp": c["warp_nodes"] / (c["rays"] / 32), "recs_per_warp": c["warp_leaf_prims"] / (c["rays"] / 32)}))
ctx.close()
