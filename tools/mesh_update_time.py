"""Cost of moving a reflection primitive (updateInstanceTransforms): grtSetMeshes + grtBuildMeshAccel per frame."""
import importlib
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
grt = importlib.import_module("gaussian-ray-tracing_b200")

ctx = grt.Context(0)
plane = grt.create_plane([0, 0, 0])
sphere = grt.create_sphere([0.45, 0.0, 1.8])
set_ms, build_ms, dev_ms = [], [], []
for k in range(12):
    sphere["transform"] = grt.scene.compose(grt.translate([0.45 + 0.01 * k, 0.0, 1.8]))
    plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0 + k), grt.scale_mat(3.0))
    t0 = time.perf_counter()
    ctx.set_meshes([plane, sphere])
    t1 = time.perf_counter()
    ctx.build_mesh_accel()
    t2 = time.perf_counter()
    set_ms.append((t1 - t0) * 1e3)
    build_ms.append((t2 - t1) * 1e3)
    dev_ms.append(ctx.build_info()["mesh_build_ms"])
ctx.validate_accel(1)
info = ctx.build_info()
print(json.dumps(dict(triangles=info["n_mesh_triangles"], set_meshes_wall_ms=round(min(set_ms[2:]), 3),
                      build_wall_ms=round(min(build_ms[2:]), 3), build_device_ms=round(min(dev_ms[2:]), 3))))
ctx.close()
