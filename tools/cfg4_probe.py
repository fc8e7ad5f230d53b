#!/usr/bin/env python
"""Step-by-step probe of configs[3] (1M Gaussians + mirror plane and sphere) with a print after every host call."""
import importlib, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
grt = importlib.import_module("gaussian-ray-tracing_b200")
from conftest import make_scene
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
W, H = (1920, 1080) if n >= 500_000 else (192, 128)
t0 = time.time()
def say(*a): print(f"[{time.time()-t0:6.1f}s]", *a, flush=True)
rec, P, _ = make_scene(grt, n, 1002, W, H, deg=3)
ctx = grt.Context(0); say("context")
ctx.set_particles(rec, 0.01); P.handle, P.d_particles = ctx.build_particle_accel(); say("particle accel", ctx.build_info()["build_ms"])
plane = grt.create_plane([0, 0, 0])
plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
sphere = grt.create_sphere([0.45, 0.0, 1.8])
d = ctx.malloc(W * H * 3); P.output_buffer = d
ctx.launch(P); ctx.sync(); say("plain launch ms", ctx.last_launch_ms())
ctx.set_meshes([plane, sphere]); say("set_meshes")
P.mesh_handle, P.d_meshes = ctx.build_mesh_accel(); say("mesh accel", ctx.build_info()["mesh_build_ms"])
for counters in (False, True):
    ctx.enable_counters(counters)
    ctx.launch(P); say("mesh launch issued, counters", counters); ctx.sync(); say("mesh launch ms", ctx.last_launch_ms())
c = ctx.counters(); say({k: c[k] for k in ("rays", "segments", "stack_overflow", "warp_per_ray_passes", "lanes_repass", "warp_dfs_fallbacks")})
