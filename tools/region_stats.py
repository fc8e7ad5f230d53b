#!/usr/bin/env python
"""Traversal counters of the bench frame per image region (background corner vs sphere centre vs ground)."""
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
regions = {"corner_bg": (0, 0, 512, 256), "centre_sphere": (832, 412, 256, 256), "bottom_ground": (704, 824, 512, 256), "full": (0, 0, W, H)}
for name, (x0, y0, w, h) in regions.items():
    ctx.enable_counters(True); ctx.counters(reset=True)
    ctx.launch_tile(P, x0, y0, w, h); ctx.sync()
    c = ctx.counters(); ctx.enable_counters(False)
    ctx.launch_tile(P, x0, y0, w, h); ctx.sync()
    ms = ctx.last_launch_ms()
    rays = c["rays"]; warps = rays / 32; d = c["dbg"]
    print(f"{name:14s} ms {ms:6.3f} rays {rays} hg/ray {c['hits_gathered']/rays:5.1f} hs/ray {c['hits_shaded']/rays:5.1f} "
          f"discovered_events/ray {d[0]/rays:5.1f} steps/insert {d[1]/max(d[0],1):4.2f} cnt@insert {d[2]/max(d[0],1):4.1f} "
          f"cand/ray {d[3]/rays:5.1f} hitpairs/ray {d[4]/rays:5.1f} spill/ray {d[5]/rays:4.1f} nodes/w {c['warp_nodes']/warps:5.1f} "
          f"recs/w {c['warp_leaf_prims']/warps:5.1f} cand_iters/w {c['warp_candidate_iters']/warps:5.1f} shade_iters/w {c['warp_consume_iters']/warps:5.1f} repass {c['lanes_repass']}")
ctx.close()
