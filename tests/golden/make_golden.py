#!/usr/bin/env python
"""Generates tests/golden/*.npz: small frames rendered by the CPU oracle (oracle/oracle.cpp, brute-force AND BVH mode
agree bit for bit, see tests/test_oracle_kat.py) for the synthetic scene generator of SURVEY 8d.

The reference ships no golden vectors and its OptiX code cannot run here, so these frames pin
(a) the oracle against silent drift (tests/test_golden.py::test_oracle_reproduces_golden, CPU) and
(b) the CUDA path on the GPU box, where /root/reference does not exist (tests/test_golden.py::test_gpu_matches_golden).

usage: python tests/golden/make_golden.py        (from the repo root; needs oracle/liboracle.so)
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CASES = {
    # name: (n, seed, width, height, deg, fisheye, mesh_type, with_meshes)
    "pinhole_deg3": (3000, 201, 64, 48, 3, False, 0, False),
    "pinhole_deg0": (3000, 202, 64, 48, 0, False, 0, False),
    "fisheye_deg3": (3000, 203, 56, 56, 3, True, 0, False),
    "mirror_deg3": (3000, 204, 64, 48, 3, False, 0, True),
    "glass_deg2": (3000, 205, 64, 48, 2, False, 2, True),
    "normal_deg1": (3000, 206, 64, 48, 1, False, 1, True),
}


def case_inputs(grt, name):
    from conftest import make_scene
    n, seed, w, h, deg, fisheye, mesh_type, with_meshes = CASES[name]
    rec, P, _ = make_scene(grt, n, seed, w, h, deg=deg, fisheye=fisheye, mesh_type=mesh_type)
    meshes = None
    if with_meshes:
        plane = grt.create_plane([0, 0, 0])
        plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
        meshes = [plane, grt.create_sphere([0.45, 0.0, 1.8])]
    return rec, P, meshes


def main():
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    from oracle import orc
    for name in CASES:
        rec, P, meshes = case_inputs(grt, name)
        S = orc.Scene(rec, 0.01)
        if meshes:
            S.set_meshes(meshes)
        r = S.render(orc.params_from_grt(P), K=7)
        out = os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".npz")
        np.savez_compressed(out, rgb8=r["rgb8"], rgb32f=r["rgb32f"].astype(np.float32), flags=r["flags"].astype(np.uint8),
                            hits_gathered=np.int64(r["stats"]["hits_gathered"]), hits_shaded=np.int64(r["stats"]["hits_shaded"]))
        print(name, r["rgb8"].shape, "flagged", int(r["flags"].astype(bool).sum()), "h_g", r["stats"]["hits_gathered"], os.path.getsize(out), "bytes")


if __name__ == "__main__":
    main()
