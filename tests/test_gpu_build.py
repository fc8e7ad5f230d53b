"""-m gpu: the on-device LBVH build (build.cu) through the C ABI."""
import numpy as np
import pytest

from conftest import make_scene

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n", [1, 2, 255, 256, 257, 4096, 4097, 100_003, 1_000_000])
def test_radix_sort_matches_numpy(ctx, n):
    rng = np.random.default_rng(n)
    keys = rng.integers(0, 2**63, size=n, dtype=np.uint64)
    if n > 10:
        keys[: n // 3] = keys[n // 3: 2 * (n // 3)][: n // 3]  # many duplicates -> stability matters
    vals = np.arange(n, dtype=np.uint32)
    k2, v2 = ctx.radix_sort(keys, vals)
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k2, keys[order])
    assert np.array_equal(v2, vals[order])  # bit-exact incl. stable order of equal keys


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 5, 9, 64, 1000, 50_000])
def test_particle_accel_valid(ctx, grt, n):
    rec, P, _ = make_scene(grt, max(n, 1), 11 + n, 64, 64)
    rec = rec[:n]
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    info = ctx.build_info()
    live = int((rec[:, 10] > 0.01).sum()) if n else 0
    assert info["n_particles"] == n
    assert info["n_live"] == live
    assert (handle != 0) == (live > 0)
    ctx.validate_accel(0)


def test_all_culled_and_duplicates(ctx, grt):
    rec, P, _ = make_scene(grt, 500, 5, 64, 64)
    dead = rec.copy()
    dead[:, 10] = 0.005  # opacity <= alpha_min: proxy scale NaN/0 in the reference -> never hit
    ctx.set_particles(dead, 0.01)
    handle, _ = ctx.build_particle_accel()
    assert handle == 0 and ctx.build_info()["n_live"] == 0
    same = np.repeat(rec[:1], 300, axis=0)  # 300 identical particles: identical Morton codes
    ctx.set_particles(same, 0.01)
    handle, _ = ctx.build_particle_accel()
    assert ctx.build_info()["n_live"] == (300 if rec[0, 10] > 0.01 else 0)
    ctx.validate_accel(0)


def test_mesh_accel_valid(ctx, grt):
    ctx.set_meshes([grt.create_plane([0, 0, 1]), grt.create_sphere([0.5, 0, 1.5])])
    handle, dm = ctx.build_mesh_accel()
    info = ctx.build_info()
    assert handle != 0 and dm
    assert info["n_mesh_triangles"] == 2 + 6 * 180 * 89 // 3
    ctx.validate_accel(1)
    ctx.set_meshes([])
    handle, dm = ctx.build_mesh_accel()
    assert handle == 0 and not dm
