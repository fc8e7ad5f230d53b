"""Dynamic particle sets: grtSetParticles* + grtBuildParticleAccel per frame reuse the context's scratch
block and accel arrays.  A rebuilt accel must be indistinguishable from a fresh context's, whatever was built before."""
import numpy as np
import pytest

from conftest import make_scene
from gpu_util import gpu_render

pytestmark = pytest.mark.gpu


def test_rebuild_sequence_matches_fresh_context(grt, ctx):
    W, H = 160, 96
    frames = {}
    # grow, shrink, grow again, empty, and back: every build reuses or regrows the arrays of the one before
    # (24 000 after 20 000: the kept node array passes the builder's size guess but is too small in the end, so the collapse
    # is repeated into scratch -- csrc/build.cu, second attempt)
    for n, seed in [(20_000, 5), (24_000, 11), (60_000, 6), (3_000, 7), (60_000, 6), (0, 8), (20_000, 5)]:
        if n == 0:
            ctx.set_particles(np.zeros((0, 59), np.float32), 0.01)
            handle, _ = ctx.build_particle_accel()
            assert handle == 0
            continue
        rec, P, _ = make_scene(grt, n, seed, W, H, deg=3)
        img = gpu_render(ctx, grt, rec, P).rgb8
        assert ctx.build_info()["n_live"] > 0.98 * n
        if (n, seed) in frames:
            assert np.array_equal(frames[(n, seed)], img)  # same scene after other builds in between: same frame
        frames[(n, seed)] = img
    fresh = grt.Context(0)
    try:
        for (n, seed), img in frames.items():
            rec, P, _ = make_scene(grt, n, seed, W, H, deg=3)
            assert np.array_equal(gpu_render(fresh, grt, rec, P).rgb8, img)
    finally:
        fresh.close()


def test_stale_handle_is_rejected_after_new_particles(grt, ctx):
    rec, P, _ = make_scene(grt, 2_000, 9, 64, 48, deg=1)
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    P.handle, P.d_particles = handle, dpart
    d_out = ctx.malloc(64 * 48 * 3)
    P.output_buffer = d_out
    ctx.launch(P)
    ctx.sync()
    ctx.set_particles(rec[:1000], 0.01)          # new particle set, accel not rebuilt yet
    with pytest.raises(grt.GrtError, match="particle accel"):
        ctx.launch(P)
    handle, dpart = ctx.build_particle_accel()
    P.handle, P.d_particles = handle, dpart
    ctx.launch(P)
    ctx.sync()
    ctx.free(d_out)


def test_rebuild_time_1m(grt, ctx):
    """Per-frame rebuild of a 1M-particle accel: the second and later builds take their temporaries from the scratch block."""
    rec = grt.activate(grt.synth_ply_properties(1_000_000, 1002))
    ctx.set_particles(rec, 0.01)
    ms = []
    import time
    wall = []
    for _ in range(4):
        t = time.perf_counter()
        ctx.build_particle_accel()
        wall.append((time.perf_counter() - t) * 1e3)
        ms.append(ctx.build_info()["build_ms"])
    ctx.validate_accel(0)
    print("1M rebuild device ms", [round(m, 2) for m in ms], "wall ms", [round(w, 2) for w in wall],
          "Gprims/s", round(1e-6 / (min(ms[1:]) * 1e-3), 3))
    # >= 1 Gprims/s would be 1.0 ms; measured 0.98 ms on a quiet B200 -- the bound leaves room for clock and
    # allocator jitter of a shared box, not for a regression to the 2 ms of the unpooled builder
    assert min(ms[1:]) < 1.5, ms
    assert min(wall[1:]) < 3.0, wall
    ctx.set_particles(rec[:0], 0.01)


def test_rebuilt_accel_renders_like_the_oracle(grt, ctx, orc):
    """A frame after grtSetParticles + rebuild is compared with the ORACLE of the new particle set (not only with a fresh
    context): moved particles, a changed count and changed opacities (some cross the cull threshold alpha_min)."""
    from gpu_util import compare
    W, H = 160, 96
    rec, P, _ = make_scene(grt, 30_000, 21, W, H, deg=3)
    gpu_render(ctx, grt, rec, P)                                   # the accel the next builds reuse
    rng = np.random.default_rng(3)
    rec2 = rec[: 25_000].copy()
    rec2[:, 0:3] += rng.normal(0, 0.05, (len(rec2), 3)).astype(np.float32)
    rec2[::7, 10] = np.float32(0.009)                              # culled now (opacity <= alpha_min)
    rec2[1::7, 10] = np.float32(0.95)
    for frame in (rec2, rec):                                      # shrink + move, then grow back
        gpu = gpu_render(ctx, grt, frame, P, counters=True)
        S = orc.Scene(frame, 0.01)
        ora = S.render(orc.params_from_grt(P), K=7)
        compare(gpu, ora, label=f"rebuilt accel, {len(frame)} particles")
        assert ctx.build_info()["n_live"] == S.live_count()
        assert gpu.counters["stack_overflow"] == 0
