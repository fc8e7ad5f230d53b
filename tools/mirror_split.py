import importlib, os, sys
import numpy as np
sys.path.insert(0, "/root/repo")
grt = importlib.import_module("gaussian-ray-tracing_b200")
W, H = 1920, 1080
rec = grt.activate(grt.synth_ply_properties(1_000_000, 1002))
ctx = grt.Context(0); ctx.set_particles(rec, 0.01)
handle, dpart = ctx.build_particle_accel()
cam = grt.default_camera(rec)
U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3)
P.handle, P.d_particles = handle, dpart
d_out = ctx.malloc(W * H * 3); P.output_buffer = d_out
plane = grt.create_plane([0, 0, 0])
plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
ctx.set_meshes([plane, grt.create_sphere([0.45, 0.0, 1.8])])
P.mesh_handle, P.d_meshes = ctx.build_mesh_accel(); P.type = 0
for name, t in [("full", (0, 0, W, H)), ("hot block r4c8", (960, 480, 120, 120)), ("sphere rim r5c3", (360, 600, 120, 120)), ("plane interior r1c10", (1200, 120, 120, 120))]:
    ctx.enable_counters(True); ctx.counters(reset=True)
    ctx.launch_tile(P, *t); ctx.sync()
    c = ctx.counters(); ctx.enable_counters(False)
    d = c["dbg"]
    print(name, "mesh cycles %.3g gauss cycles %.3g ratio %.2f segs/ray %.2f nodes/w %.0f recs/w %.0f" % (d[6], d[7], d[6] / max(d[7], 1), c["segments"] / c["rays"], c["warp_nodes"] / (c["rays"] / 32), c["warp_leaf_prims"] / (c["rays"] / 32)))
