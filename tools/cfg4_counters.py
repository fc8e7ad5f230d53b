"""configs[3] (mirror plane + sphere) with the counting kernel: prints the traversal counters."""
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
plane = grt.create_plane([0, 0, 0])
plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
sphere = grt.create_sphere([0.45, 0.0, 1.8])
ctx.set_meshes([plane, sphere])
mh, dm = ctx.build_mesh_accel()
cam = grt.default_camera(rec)
U, V, Wv = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(W) / np.float32(H)))
P = grt.make_params(W, H, cam["eye"], U, V, Wv, sh_degree_max=3)
P.handle, P.d_particles, P.mesh_handle, P.d_meshes, P.type = handle, dpart, mh, dm, 0
d_out = ctx.malloc(W * H * 3)
P.output_buffer = d_out
tiles = [(0, 0, W, H)]
if len(sys.argv) > 1 and sys.argv[1] == "--grid":
    tiles = [(x, y, 240, 135) for y in range(0, H, 135) for x in range(0, W, 240)]
for t in tiles:
    ctx.enable_counters(True); ctx.counters(reset=True)
    ctx.launch_tile(P, *t); ctx.sync()
    c = ctx.counters(); ctx.enable_counters(False)
    if len(tiles) == 1 or c["stack_overflow"] or c["order_violations"]:
        print(t, json.dumps({k: v for k, v in c.items() if k != "dbg"}))
ctx.launch(P); ctx.sync()
print("ms", ctx.last_launch_ms())
ctx.close()
