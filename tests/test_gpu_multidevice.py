"""Contexts are bound to ONE GPU each and several may live in one process (include/grt.h: grtCreate(device, ...)).
The dynamic-shared-memory opt-in of the trace kernel is a per-device attribute and every entry point has to make its context's
device current (a process-wide flag and a missing cudaSetDevice broke the second GPU of a process)."""
import ctypes as C

import numpy as np
import pytest

from conftest import make_scene
from gpu_util import gpu_render

pytestmark = pytest.mark.gpu


def _device_count(grt):
    import torch
    return torch.cuda.device_count()


def test_two_contexts_on_one_device_do_not_share_state(grt, ctx):
    """Interleaved use of two contexts on device 0: separate scenes, handles, tile counters and debug outputs."""
    other = grt.Context(0)
    try:
        rec_a, P_a, _ = make_scene(grt, 4_000, 31, 96, 64, deg=3)
        rec_b, P_b, _ = make_scene(grt, 6_000, 32, 96, 64, deg=1)
        a1 = gpu_render(ctx, grt, rec_a, P_a).rgb8
        b1 = gpu_render(other, grt, rec_b, P_b).rgb8
        a2 = gpu_render(ctx, grt, rec_a, P_a).rgb8
        assert np.array_equal(a1, a2) and not np.array_equal(a1, b1)
        # a handle of one context is rejected by the other
        ctx.set_particles(rec_a, 0.01)
        P_a.handle, P_a.d_particles = ctx.build_particle_accel()
        d = other.malloc(96 * 64 * 3)
        P_a.output_buffer = d
        with pytest.raises(grt.GrtError):
            other.launch(P_a)
        other.free(d)
        P_a.output_buffer = None
    finally:
        other.close()


def test_contexts_on_two_devices_in_one_process(grt, ctx):
    """Same frame from device 0 and device 1 of ONE process, with the caller's current device left on the other GPU."""
    if _device_count(grt) < 2:
        pytest.skip("needs 2 GPUs in one box (run under gpurun --gpus 2; log kept in profiles/)")
    import torch
    rec, P, _ = make_scene(grt, 20_000, 33, 192, 128, deg=3)
    ref = gpu_render(ctx, grt, rec, P).rgb8
    torch.cuda.set_device(0)
    dev1 = grt.Context(1)
    try:
        torch.cuda.set_device(0)                       # the caller's current device is NOT the context's
        img1 = gpu_render(dev1, grt, rec, P).rgb8
        assert np.array_equal(img1, ref)
        torch.cuda.set_device(1)
        img0 = gpu_render(ctx, grt, rec, P).rgb8        # and the other way round
        assert np.array_equal(img0, ref)
        out = dev1.pinned_empty((128, 192, 3))
        dev1.set_particles(rec, 0.01)
        P.handle, P.d_particles = dev1.build_particle_accel()
        torch.cuda.set_device(0)
        dev1.render_host(P, out)                        # allocates its device frame on device 1
        assert np.array_equal(out, ref)
    finally:
        dev1.close()
        torch.cuda.set_device(0)


@pytest.mark.parametrize("mode", ["bands", "cameras"])
def test_nccl_gather_through_the_c_abi(grt, tmp_path, mode):
    """One process per GPU, NO torch.distributed: row bands of one frame (grtLaunchRows + grtGatherBands) and a camera batch
    (grtLaunch + grtGatherFrames) over libgrt's own NCCL exchange reproduce what one GPU renders alone, bit for bit."""
    import os
    import subprocess
    import sys
    world = min(_device_count(grt), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs in one box (run under gpurun --gpus 2; log kept in profiles/)")
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mgpu_worker.py")
    id_file, out = str(tmp_path / "nccl_id"), str(tmp_path / "out.npz")
    procs = [subprocess.Popen([sys.executable, worker, str(r), str(world), id_file, out, mode], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
             for r in range(world)]
    logs = []
    for p in procs:
        try:
            logs.append(p.communicate(timeout=300)[0])
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
    assert all(p.returncode == 0 for p in procs), "\n".join(logs)
    z = np.load(out)
    assert z["got"].shape == z["want"].shape and z["want"].max() > 0
    assert np.array_equal(z["got"], z["want"])
