"""Device time of grtBuildParticleAccel (proxy boxes + LBVH + record gather) for rebuilds of a resident particle set."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import importlib

grt = importlib.import_module("gaussian-ray-tracing_b200")


def main():
    out = []
    ctx = grt.Context(0)
    sizes = [(100_000, 1001), (1_000_000, 1002), (6_000_000, 1003)]
    if len(sys.argv) > 1:
        sel = [a for a in sys.argv[1:] if a.isdigit()]
        sizes = [s for s in sizes if str(s[0]) in sel] if sel else sizes
    reps = 3 if "--short" in sys.argv else 8
    for n, seed in sizes:
        rec = grt.activate(grt.synth_ply_properties(n, seed))
        ctx.set_particles(rec, 0.01)
        ms, wall = [], []
        for _ in range(reps):
            t = time.perf_counter()
            ctx.build_particle_accel()
            wall.append((time.perf_counter() - t) * 1e3)
            ms.append(ctx.build_info()["build_ms"])
        ctx.validate_accel(0)
        info = ctx.build_info()
        best = min(ms[1:])
        out.append(dict(n=n, n_live=info["n_live"], wide_nodes=info["n_wide_nodes"], levels=info["wide_depth_levels"],
                        first_ms=round(ms[0], 3), rebuild_ms=round(best, 3), rebuild_wall_ms=round(min(wall[1:]), 3),
                        gprims_per_s=round(n / best / 1e6, 3)))
        print(json.dumps(out[-1]), flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
