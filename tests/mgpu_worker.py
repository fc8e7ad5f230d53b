"""Worker of the multi-GPU C-ABI tests (one process per GPU, no torch.distributed, no gloo): usage
    mgpu_worker.py RANK WORLD ID_FILE OUT_NPZ MODE        MODE = bands | cameras
Every rank builds the same scene on ITS GPU, traces its share through the C ABI and takes part in the NCCL gather of libgrt
(grtCommInit / grtGatherBands / grtGatherFrames).  Rank 0 also traces the whole work alone and writes both results."""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world, id_file, out_npz, mode = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], sys.argv[5]
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    from conftest import make_scene
    W, H = 320, 200
    rec, P, cam = make_scene(grt, 30_000, 77, W, H, deg=3)
    ctx = grt.Context(rank)
    if rank == 0:
        uid = grt.Context.comm_unique_id()
        with open(id_file + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(id_file + ".tmp", id_file)
    else:
        t0 = time.time()
        while not os.path.exists(id_file):
            if time.time() - t0 > 120:
                raise SystemExit("rank 0 never wrote the NCCL id")
            time.sleep(0.05)
        uid = open(id_file, "rb").read()
    ctx.comm_init(world, rank, uid)
    ctx.set_particles(rec, 0.01)
    P.handle, P.d_particles = ctx.build_particle_accel()
    n = W * H * 3
    if mode == "bands":
        d = ctx.malloc(n)
        ctx.memset(d, 0, n)
        P.output_buffer = d
        ctx.launch_rows(P, 8, world, rank)
        ctx.gather_bands(d, W, H, 8, 0)
        ctx.sync()
        if rank == 0:
            got = ctx.to_host(d, (H, W, 3), np.uint8)
            ctx.memset(d, 0, n)
            ctx.launch(P)
            ctx.sync()
            np.savez(out_npz, got=got, want=ctx.to_host(d, (H, W, 3), np.uint8))
    else:  # camera k -> rank k mod world, 2 cameras per rank, frames gathered camera by camera
        def params_for(k):
            th = 2 * np.pi * k / (2 * world)
            eye = np.array([3 * np.sin(th), 0.3, 3 * np.cos(th)], np.float32)
            U, V, Wv = grt.camera_uvw(eye, cam["lookat"], cam["up"], 60.0, float(np.float32(W) / np.float32(H)))
            Q = grt.make_params(W, H, eye, U, V, Wv, sh_degree_max=3)
            Q.handle, Q.d_particles = P.handle, P.d_particles
            return Q
        d = ctx.malloc(n)
        d_all = ctx.malloc(n * world) if rank == 0 else None
        got = []
        for rnd in range(2):
            Q = params_for(rnd * world + rank)
            Q.output_buffer = d
            ctx.launch(Q)
            ctx.gather_frames(d, n, d_all, 0)
            ctx.sync()
            if rank == 0:
                got.append(ctx.to_host(d_all, (world, H, W, 3), np.uint8))
        if rank == 0:
            want = []
            for k in range(2 * world):
                Q = params_for(k)
                Q.output_buffer = d
                ctx.launch(Q)
                ctx.sync()
                want.append(ctx.to_host(d, (H, W, 3), np.uint8))
            np.savez(out_npz, got=np.concatenate(got), want=np.stack(want))
    ctx.comm_destroy()
    ctx.close()
    print(f"rank {rank} ok", flush=True)


if __name__ == "__main__":
    main()
