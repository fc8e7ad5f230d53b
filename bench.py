#!/usr/bin/env python
"""bench.py -- headline measurement: Mrays/s at 1M Gaussians, 1920x1080 pinhole, SH degree 3 (BASELINE.json configs[1]).

One "step" = one pass of the hot path over one batch of synthetic input = one full 1080p frame per GPU.
N > 1 (torchrun, one rank per GPU): camera-batch partition -- rank r traces camera r of an N-camera ring
(scene + BVH replicated per GPU); the frames are gathered to rank 0 by libgrt's own NCCL exchange
(grtGatherFrames) on a side stream, double-buffered, so that the gather of step k runs under the kernel of step k+1;
everything is inside the timed span ("weak" scaling: per-GPU work is fixed).

The same run then measures, outside the headline's timed region, and adds to the JSON line:
  "strong"  -- ONE 1080p frame (camera 0) split into interleaved 8-row bands over the N ranks, one grtLaunchRows per rank +
               grtGatherBands: ms/frame and the speed-up over one GPU;
  "config5" -- BASELINE configs[4]: 6M Gaussians, 3840x2160, the 64-camera ring, camera k -> rank k mod N, gathers overlapped;
  config.rebuild_ms / rebuild_gprims_s -- an event-timed rebuild of the same 1M accel.

JSON keys follow the driver contract; `roofline` and `cpu_baseline` are described in DESIGN.md section 6.
`--impl reference` times the CPU oracle (the reference has no CPU path and its OptiX build is unavailable on
this box: BASELINE.md section 2) on all host cores on a bounded sample of the same workload.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = dict(n=1_000_000, seed=1002, width=1920, height=1080, deg=3)
SAMPLE_ROWS = 24  # CPU sample: a centred band of this many full-width rows of the same frame


def camera_for_rank(grt, rec, rank, n_ranks, width, height, deg, fisheye=False):
    cam = grt.default_camera(rec)
    theta = 2.0 * np.pi * rank / max(n_ranks, 1)
    eye = np.array([3.0 * np.sin(theta), 0.0, 3.0 * np.cos(theta)], np.float32)  # rank 0 = reference default eye (0,0,3)
    U, V, W = grt.camera_uvw(eye, cam["lookat"], cam["up"], cam["fov_y"], float(np.float32(width) / np.float32(height)))
    return grt.make_params(width, height, eye, U, V, W, sh_degree_max=deg, fisheye=fisheye)


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(n)
            except Exception:
                pass
            time.sleep(0.1)

    def result(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def run_reference(args, rank):
    """CPU arm: the oracle (kind 'port'), all host threads, bounded sample per step."""
    if rank != 0:
        return
    grt = importlib.import_module("gaussian-ray-tracing_b200")  # harness only: scene generator / camera
    from oracle import orc
    w = WORKLOAD
    rec = grt.activate(grt.synth_ply_properties(w["n"], w["seed"]))
    P = camera_for_rank(grt, rec, 0, 1, w["width"], w["height"], w["deg"])
    S = orc.Scene(rec, 0.01)
    y0 = (w["height"] - SAMPLE_ROWS) // 2
    rect = (0, y0, w["width"], SAMPLE_ROWS)
    op = orc.params_from_grt(P)
    secs, threads = [], 0
    for i in range(args.warmup + args.steps):
        r = S.render(op, K=7, rect=rect, want_float=False, want_flags=False)
        if i >= args.warmup:
            secs.append(r["stats"]["seconds"])
        threads = r["stats"]["threads"]
    rays = w["width"] * SAMPLE_ROWS
    val = rays * len(secs) / sum(secs) / 1e6
    sample = f"{SAMPLE_ROWS} centred full-width rows ({rays} rays) of the 1920x1080 frame per step, K=7"
    print(json.dumps({
        "impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sum(secs) / len(secs), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "1M synthetic Gaussians (seed 1002), 1920x1080 pinhole, SH deg 3", "sample": sample},
        "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


def sha256_of(path):
    import hashlib
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def ring_camera(grt, lookat, up, k, n, width, height, deg, y=0.5):
    """camera k of config 5's ring: eye = (3 sin th, 0.5, 3 cos th), th = 2 pi k / n"""
    th = 2.0 * np.pi * k / n
    eye = np.array([3.0 * np.sin(th), y, 3.0 * np.cos(th)], np.float32)
    U, V, W = grt.camera_uvw(eye, lookat, up, 60.0, float(np.float32(width) / np.float32(height)))
    return grt.make_params(width, height, eye, U, V, W, sh_degree_max=deg)


def config4_meshes(grt):
    plane = grt.create_plane([0, 0, 0])
    plane["transform"] = grt.scene.compose(grt.translate([-0.35, 0.0, 1.6]), grt.rotate_y(35.0), grt.scale_mat(3.0))
    return [plane, grt.create_sphere([0.45, 0.0, 1.8])]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="grt", choices=["grt", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fisheye", action="store_true", help="same as --config 3")
    ap.add_argument("--config", type=int, default=2, choices=[2, 3, 4],
                    help="builder's switch: BASELINE configs[1] pinhole (default, the headline), [2] fisheye, [3] mirror plane + sphere")
    ap.add_argument("--no-extras", action="store_true", help="skip the strong-scaling and config 5 blocks")
    ap.add_argument("--config5-cameras", type=int, default=64)
    ap.add_argument("--strong-frames", type=int, default=10)
    args = ap.parse_args()
    if args.fisheye:
        args.config = 3
    args.warmup = max(args.warmup, 3) if args.impl == "grt" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    ctx = grt.Context(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))   # plumbing: barrier, max over ranks, the NCCL id
        uid = [grt.Context.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.comm_init(world, rank, uid[0])                                             # the data path: libgrt's own communicator
    w = WORKLOAD
    W, H = w["width"], w["height"]
    rays = W * H
    nbytes = rays * 3
    rec = grt.activate(grt.synth_ply_properties(w["n"], w["seed"]))
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    info = ctx.build_info()
    rebuild_ms = []
    for _ in range(3):                       # dynamic particle sets: the same set rebuilt (arrays and scratch block reused)
        handle, dpart = ctx.build_particle_accel()
        rebuild_ms.append(float(ctx.build_info()["build_ms"]))
    fisheye = args.config == 3
    P = camera_for_rank(grt, rec, rank, world, W, H, w["deg"], fisheye=fisheye)
    P.handle, P.d_particles = handle, dpart
    if args.config == 4:
        ctx.set_meshes(config4_meshes(grt))
        P.mesh_handle, P.d_meshes = ctx.build_mesh_accel()
    frames = [torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda") for _ in range(2)]
    gathered = [torch.empty((world, H, W, 3), dtype=torch.uint8, device="cuda") for _ in range(2)] if (world > 1 and rank == 0) else [None, None]
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    main_s = torch.cuda.current_stream()
    side_s = torch.cuda.Stream()

    # one counted pass (separate kernel variant, outside every timed region) for h_g / h_s
    P.output_buffer = frames[0].data_ptr()
    ctx.enable_counters(True)
    ctx.counters(reset=True)
    ctx.launch(P)
    torch.cuda.synchronize()
    cnt = ctx.counters()
    ctx.enable_counters(False)
    assert cnt["stack_overflow"] == 0, cnt

    def run_steps(n_steps, timed):
        """n_steps frames; the gather of frame k (side stream) overlaps flush + kernel of frame k+1 (main stream).
        Returns (span_ms, flush_ms, kernel_ms list): span = first kernel start .. last gather end."""
        E = lambda: torch.cuda.Event(enable_timing=True)
        evs = [(E(), E(), E(), E(), E()) for _ in range(n_steps)]   # flush start, kernel start, kernel end, gather end, flush end
        for k, (f0, a, b, g, f1) in enumerate(evs):
            buf = k & 1
            f0.record(main_s)
            flush.zero_()                       # L2 flush between timed iterations
            f1.record(main_s)
            if k >= 2:
                main_s.wait_event(evs[k - 2][3])  # frame buffer `buf` is free again once its gather is through
            a.record(main_s)
            P.output_buffer = frames[buf].data_ptr()
            ctx.launch(P)                       # default stream == torch's current stream
            b.record(main_s)
            if world > 1:
                side_s.wait_event(b)
                ctx.gather_frames(frames[buf].data_ptr(), nbytes, gathered[buf].data_ptr() if rank == 0 else None, 0, stream=side_s.cuda_stream)
            g.record(side_s if world > 1 else main_s)
        main_s.wait_stream(side_s)
        end = E()
        end.record(main_s)
        torch.cuda.synchronize()
        if not timed:
            return None
        span = evs[0][0].elapsed_time(end)
        fl = sum(f0.elapsed_time(f1) for f0, _, _, _, f1 in evs)  # the flush writes only: a wait for a frame buffer still being gathered stays in the span
        return span, fl, [a.elapsed_time(b) for _, a, b, _, _ in evs]

    run_steps(args.warmup, False)
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    launches_before = ctx.counters()["kernel_launches"]
    torch.cuda.synchronize()
    span_ms, flush_ms, kernel_ms = run_steps(args.steps, True)
    if world > 1:
        dist.barrier()
    sampler.stop_flag = True
    sampler.join(timeout=3)
    total_ms = span_ms - flush_ms           # the flush writes sit between the frames on the same stream; they are not part of a step
    kernel_ms_max_rank = float(np.mean(kernel_ms))
    if world > 1:
        t = torch.tensor([total_ms, kernel_ms_max_rank], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, kernel_ms_max_rank = float(t[0].item()), float(t[1].item())
    launches = ctx.counters()["kernel_launches"] - launches_before
    value = world * rays * args.steps / (total_ms * 1e-3) / 1e6
    last = (args.steps - 1) & 1
    if world > 1 and rank == 0:
        assert torch.equal(gathered[last][0], frames[last]), "gathered frame of rank 0 differs from its own frame"
        assert int(gathered[last][world - 1].max()) > 0, "the last rank's frame did not arrive"

    # e2e through the C ABI with HOST buffers: Params from host memory, frame copied back to pinned host memory; L2 flushed before
    # every step like in the device-timed loop (outside the per-step wall-clock bracket)
    host = ctx.pinned_empty((H, W, 3))
    Pe = camera_for_rank(grt, rec, rank, world, W, H, w["deg"], fisheye=fisheye)
    Pe.handle, Pe.d_particles, Pe.mesh_handle, Pe.d_meshes = P.handle, P.d_particles, P.mesh_handle, P.d_meshes
    for _ in range(2):
        ctx.render_host(Pe, host)
    if world > 1:
        dist.barrier()
    e2e_s = 0.0
    for _ in range(args.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        ctx.render_host(Pe, host)
        e2e_s += time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = world * rays * args.steps / e2e_s / 1e6
    assert np.array_equal(host, frames[last].cpu().numpy()), "e2e frame differs from the device-resident frame"

    extras = {}
    if not args.no_extras and args.config == 2:
        extras["strong"] = strong_scaling_block(args, grt, ctx, torch, dist, rec, handle, dpart, flush, rank, world, W, H, w["deg"], float(np.mean(kernel_ms)))

    line = None
    if rank == 0:
        peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(peaks_path):
            peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
        else:
            peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
        hg, hs = cnt["hits_gathered"] / cnt["rays"], cnt["hits_shaded"] / cnt["rays"]
        bytes_per_ray = 44.0 * hg + 12.0 * (w["deg"] + 1) ** 2 * hs + 3.0
        k_ms = float(np.mean(kernel_ms))
        achieved = rays * bytes_per_ray / (k_ms * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum of the same kernel from the committed ncu capture of THIS kernel source:
        # refused (null) when csrc/trace.cu has changed since the capture
        traffic, traffic_src, issue_pct = None, None, None
        ncu_path = os.path.join(ROOT, "profiles", NCU_SUMMARY[args.config])
        src_sha = sha256_of(os.path.join(ROOT, "gaussian-ray-tracing_b200", "csrc", "trace.cu"))
        if os.path.exists(ncu_path):
            nj = json.load(open(ncu_path))
            if nj.get("trace_cu_sha256") == src_sha:
                traffic, issue_pct = nj.get("dram_bytes_per_launch"), nj.get("issue_active_pct")
                traffic_src = f"profiles/{NCU_SUMMARY[args.config]} (ncu --set full, captured at git {nj.get('git_head', '?')}, same csrc/trace.cu)"
            else:
                traffic_src = f"stale: profiles/{NCU_SUMMARY[args.config]} was captured for another csrc/trace.cu"
        names = {2: "pinhole", 3: "fisheye", 4: "pinhole + mirror plane and sphere (MeshType MIRROR)"}
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": "1M synthetic Gaussians (seed 1002), 1920x1080 %s, SH deg 3, one camera per GPU" % names[args.config],
                       "n_live": int(info["n_live"]), "wide_nodes": int(info["n_wide_nodes"]), "bvh_build_ms": float(info["build_ms"]),
                       "rebuild_ms": min(rebuild_ms), "rebuild_gprims_s": w["n"] / (min(rebuild_ms) * 1e-3) / 1e9,
                       "l2": "256 MB flush write before every timed frame (device-timed loop and e2e loop alike); scene arrays 256 MB + BVH > 126 MB L2",
                       "partition": ("scene+BVH replicated per GPU, one camera per rank; frames gathered to rank 0 by libgrt (grouped ncclSend/ncclRecv) "
                                     "on a side stream under the next frame's kernel") if world > 1 else "single GPU",
                       "timing": "span from the first frame's kernel start to the last gather's end, minus the L2-flush writes between frames; max over ranks",
                       "hits_gathered_per_ray": hg, "hits_shaded_per_ray": hs, "rounds_per_ray": cnt["rounds"] / cnt["rays"]},
            "kernel_ms": k_ms, "kernel_ms_max_rank": kernel_ms_max_rank,   # rank 0's camera / the most expensive camera of the ring
            "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": 144, "d2h_bytes_per_step": rays * 3},
            "gpu_launches": int(launches),
            "clocks": sampler.result(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": traffic_src,
                         "dram_frac": (traffic / (k_ms * 1e-3) / 1e9 / peak) if traffic else None,
                         "limiter": "instruction issue + dependent-instruction latency, not HBM (DESIGN.md section 5)",
                         "issue_slot_utilisation_pct": issue_pct,
                         "algorithmic_bytes_per_launch": rays * bytes_per_ray,
                         "peak_source": peak_src, "algorithmic_bytes_per_ray": bytes_per_ray,
                         "kernel": "trace_kernel<false,%s> (STATS off%s)" % ("true" if args.config == 4 else "false", "" if args.config == 4 else ", no-mesh specialisation")},
        }
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            from oracle import orc  # checker / baseline only
            S = orc.Scene(rec, 0.01)
            if args.config == 4:
                S.set_meshes(config4_meshes(grt))
            y0 = (H - SAMPLE_ROWS) // 2
            r = S.render(orc.params_from_grt(P), K=7, rect=(0, y0, W, SAMPLE_ROWS), want_float=False, want_flags=False)
            band = r["rgb8"][y0:y0 + SAMPLE_ROWS]
            gpu_band = frames[last][y0:y0 + SAMPLE_ROWS].cpu().numpy()
            line["cpu_baseline"] = {
                "value": W * SAMPLE_ROWS / r["stats"]["seconds"] / 1e6, "unit": "Mrays/s", "cores": int(r["stats"]["threads"]), "kind": "port",
                "sample": f"{SAMPLE_ROWS} centred full-width rows ({W * SAMPLE_ROWS} rays) of the same frame, K=7",
                "max_lsb_vs_gpu": int(np.abs(band.astype(int) - gpu_band.astype(int)).max()),
                "hits_gathered_per_ray": r["stats"]["hits_gathered"] / r["stats"]["rays"],
            }
            del S
    del rec
    if not args.no_extras and args.config == 2:
        c5 = config5_block(args, grt, ctx, torch, dist, rank, world, main_s, side_s)
        if rank == 0:
            line["config5"] = c5
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        ctx.comm_destroy()
        dist.destroy_process_group()


NCU_SUMMARY = {2: "r2_trace_kernel_ncu.json", 3: "r2_trace_kernel_fisheye_ncu.json", 4: "r2_mirrors_kernel_ncu.json"}


def strong_scaling_block(args, grt, ctx, torch, dist, rec, handle, dpart, flush, rank, world, W, H, deg, single_gpu_ms):
    """ONE frame (camera 0 = the reference default camera) over all ranks: interleaved 8-row bands, one grtLaunchRows per rank, then
    grtGatherBands into rank 0's frame.  Latency of the frame = kernel + gather, max over ranks."""
    P0 = camera_for_rank(grt, rec, 0, world, W, H, deg)
    P0.handle, P0.d_particles = handle, dpart
    frame = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda")
    P0.output_buffer = frame.data_ptr()
    E = lambda: torch.cuda.Event(enable_timing=True)

    def one():
        a, b, c = E(), E(), E()
        flush.zero_()
        a.record()
        if world > 1:
            ctx.launch_rows(P0, 8, world, rank)
        else:
            ctx.launch(P0)
        b.record()
        if world > 1:
            ctx.gather_bands(frame.data_ptr(), W, H, 8, 0)
        c.record()
        return a, b, c
    for _ in range(3):
        one()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    evs = [one() for _ in range(args.strong_frames)]
    torch.cuda.synchronize()
    frame_ms = float(np.mean([a.elapsed_time(c) for a, _, c in evs]))
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b, _ in evs]))
    identical = None
    if world > 1:
        t = torch.tensor([frame_ms, kern_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        frame_ms, kern_ms = float(t[0].item()), float(t[1].item())
        s1 = torch.tensor([single_gpu_ms], device="cuda")
        dist.broadcast(s1, src=0)            # rank 0 traced camera 0 alone in the headline loop
        single_gpu_ms = float(s1.item())
        if rank == 0:
            ref = torch.zeros_like(frame)
            P0.output_buffer = ref.data_ptr()
            ctx.launch(P0)
            torch.cuda.synchronize()
            identical = bool(torch.equal(ref, frame))
    return {"workload": "ONE 1920x1080 pinhole frame of the 1M scene (camera 0) split into interleaved 8-row bands over the ranks",
            "ms_per_frame": frame_ms, "kernel_ms_max_rank": kern_ms, "gather_ms": frame_ms - kern_ms, "mrays_s": W * H / (frame_ms * 1e-3) / 1e6,
            "single_gpu_ms": single_gpu_ms, "speedup_vs_1gpu": single_gpu_ms / frame_ms, "frames": args.strong_frames,
            "bit_identical_to_single_gpu_frame": identical, "launches_per_rank_per_frame": 1}


def config5_block(args, grt, ctx, torch, dist, rank, world, main_s, side_s):
    """BASELINE configs[4]: 6M Gaussians (seed 1003), 3840x2160, 64-camera ring, camera k -> rank k mod N; the gather of camera k
    (24.9 MB) runs on the side stream under the render of this rank's next camera."""
    n, w5, h5, deg, ncam = 6_000_000, 3840, 2160, 3, args.config5_cameras
    t_gen = time.perf_counter()
    rec = grt.activate(grt.synth_ply_properties(n, 1003))
    t_gen = time.perf_counter() - t_gen
    ctx.set_meshes([])
    ctx.build_mesh_accel()
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    info = ctx.build_info()
    cam = grt.default_camera(rec)
    del rec
    mine = list(range(rank, ncam, world))
    nb = w5 * h5 * 3
    frames = [torch.zeros((h5, w5, 3), dtype=torch.uint8, device="cuda") for _ in range(2)]
    recv = [torch.empty((world, h5, w5, 3), dtype=torch.uint8, device="cuda") for _ in range(2)] if (world > 1 and rank == 0) else [None, None]
    params = []
    for k in mine:
        Q = ring_camera(grt, cam["lookat"], cam["up"], k, 64, w5, h5, deg)
        Q.handle, Q.d_particles = handle, dpart
        params.append(Q)
    E = lambda: torch.cuda.Event(enable_timing=True)

    def sweep(ps):
        evs = []
        for i, Q in enumerate(ps):
            a, b, g = E(), E(), E()
            if i >= 2:
                main_s.wait_event(evs[i - 2][2])
            a.record(main_s)
            Q.output_buffer = frames[i & 1].data_ptr()
            ctx.launch(Q)
            b.record(main_s)
            if world > 1:
                side_s.wait_event(b)
                ctx.gather_frames(frames[i & 1].data_ptr(), nb, recv[i & 1].data_ptr() if rank == 0 else None, 0, stream=side_s.cuda_stream)
            g.record(side_s if world > 1 else main_s)
            evs.append((a, b, g))
        main_s.wait_stream(side_s)
        end = E()
        end.record(main_s)
        torch.cuda.synchronize()
        return evs[0][0].elapsed_time(end), [a.elapsed_time(b) for a, b, _ in evs]
    sweep(params[:2])                      # warm-up (also sizes NCCL's buffers)
    if world > 1:
        dist.barrier()
    span, kms = sweep(params)
    if world > 1:
        t = torch.tensor([span], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        span = float(t.item())
    return {"workload": f"6M synthetic Gaussians (seed 1003), 3840x2160 pinhole, SH deg 3, {ncam} cameras of the 64-camera ring, camera k -> rank k mod N",
            "cameras": ncam, "ms_total": span, "ms_per_camera": span / ncam, "kernel_ms_per_camera_rank0": float(np.mean(kms)),
            "mrays_s": ncam * w5 * h5 / (span * 1e-3) / 1e6, "n_live": int(info["n_live"]), "bvh_build_ms": float(info["build_ms"]),
            "scene_generation_s": t_gen, "l2": "no flush: 3.3 GB of scene + BVH per GPU, every camera looks from another side"}


if __name__ == "__main__":
    main()
