import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def grt():
    return importlib.import_module("gaussian-ray-tracing_b200")


@pytest.fixture(scope="session")
def orc():
    from oracle import orc as _orc
    _orc.lib()
    return _orc


@pytest.fixture(scope="session")
def ctx(grt):
    c = grt.Context(0)
    yield c
    c.close()


def make_scene(grt, n, seed, width, height, *, deg=3, fisheye=False, mesh_type=0):
    """records, GrtParams (handles unset) for the synthetic scene with the reference's default camera."""
    rec = grt.activate(grt.synth_ply_properties(n, seed))
    cam = grt.default_camera(rec)
    U, V, W = grt.camera_uvw(cam["eye"], cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(width) / np.float32(height)))
    P = grt.make_params(width, height, cam["eye"], U, V, W, sh_degree_max=deg, fisheye=fisheye, mesh_type=mesh_type)
    return rec, P, cam


def psnr(a, b):
    mse = float(np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2))
    return 99.0 if mse == 0 else 10.0 * np.log10(1.0 / mse)
