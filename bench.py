#!/usr/bin/env python
"""bench.py -- headline measurement: Mrays/s at 1M Gaussians, 1920x1080 pinhole, SH degree 3 (BASELINE.json configs[1]).

One "step" = one pass of the hot path over one batch of synthetic input = one full 1080p frame per GPU.
N > 1 (torchrun, one rank per GPU): camera-batch partition -- rank r traces camera r of an N-camera ring
(scene + BVH replicated per GPU), then the frames are gathered to rank 0 over NCCL inside the
timed region ("weak" scaling: per-GPU work is fixed).

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="grt", choices=["grt", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--fisheye", action="store_true", help="BASELINE configs[2] instead of configs[1]")
    args = ap.parse_args()
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
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    grt = importlib.import_module("gaussian-ray-tracing_b200")
    w = WORKLOAD
    W, H = w["width"], w["height"]
    rec = grt.activate(grt.synth_ply_properties(w["n"], w["seed"]))
    ctx = grt.Context(local_rank)
    ctx.set_particles(rec, 0.01)
    handle, dpart = ctx.build_particle_accel()
    info = ctx.build_info()
    P = camera_for_rank(grt, rec, rank, world, W, H, w["deg"], fisheye=args.fisheye)
    P.handle, P.d_particles = handle, dpart
    frame = torch.zeros((H, W, 3), dtype=torch.uint8, device="cuda")
    P.output_buffer = frame.data_ptr()
    part = importlib.import_module("gaussian-ray-tracing_b200.partition")
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    rays = W * H

    def step():
        ctx.launch(P)  # default stream == torch's current stream
        if world > 1:
            part.gather_frames(frame, rank, world)

    # one counted pass (separate kernel variant, outside every timed region) for h_g / h_s
    ctx.enable_counters(True)
    ctx.counters(reset=True)
    ctx.launch(P)
    torch.cuda.synchronize()
    cnt = ctx.counters()
    ctx.enable_counters(False)
    assert cnt["stack_overflow"] == 0, cnt
    launches0 = ctx.counters()["kernel_launches"]

    for _ in range(args.warmup):
        flush.zero_()
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    launches_before = ctx.counters()["kernel_launches"]
    torch.cuda.synchronize()
    for a, b, c in ev:
        flush.zero_()          # L2 flush between timed iterations (not inside the timed bracket)
        a.record()
        ctx.launch(P)
        b.record()
        if world > 1:
            part.gather_frames(frame, rank, world)  # camera-batch partition: the only exchange is the framebuffer gather
        c.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler.stop_flag = True
    sampler.join(timeout=3)
    kernel_ms = [a.elapsed_time(b) for a, b, _ in ev]
    step_ms = [a.elapsed_time(c) for a, _, c in ev]
    total_ms = float(sum(step_ms))
    if world > 1:
        t = torch.tensor([total_ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    launches = ctx.counters()["kernel_launches"] - launches_before
    value = world * rays * args.steps / (total_ms * 1e-3) / 1e6

    # e2e through the C ABI with HOST buffers: Params from host memory, frame copied back to pinned host memory
    host = ctx.pinned_empty((H, W, 3))
    Pe = camera_for_rank(grt, rec, rank, world, W, H, w["deg"], fisheye=args.fisheye)
    Pe.handle, Pe.d_particles = handle, dpart
    for _ in range(2):
        ctx.render_host(Pe, host)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ctx.render_host(Pe, host)
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e = world * rays * args.steps / e2e_s / 1e6
    assert np.array_equal(host, frame.cpu().numpy()), "e2e frame differs from the device-resident frame"

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
        traffic = None  # dram__bytes_read.sum + dram__bytes_write.sum of trace_kernel from the committed ncu capture
        ncu_path = os.path.join(ROOT, "profiles", "r1_v12_trace_kernel_ncu.json")
        if os.path.exists(ncu_path) and not args.fisheye:
            traffic = json.load(open(ncu_path)).get("dram_bytes_per_launch")
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": ("1M synthetic Gaussians (seed 1002), 1920x1080 %s, SH deg 3, one camera per GPU"
                                    % ("fisheye" if args.fisheye else "pinhole")),
                       "n_live": int(info["n_live"]), "wide_nodes": int(info["n_wide_nodes"]), "bvh_build_ms": float(info["build_ms"]),
                       "l2": "256 MB flush write between timed iterations; scene arrays 256 MB + BVH > 126 MB L2",
                       "partition": "scene+BVH replicated per GPU, one camera per rank, frames gathered to rank 0 (NCCL)" if world > 1 else "single GPU",
                       "hits_gathered_per_ray": hg, "hits_shaded_per_ray": hs, "rounds_per_ray": cnt["rounds"] / cnt["rays"]},
            "kernel_ms": k_ms,
            "e2e": {"value": e2e, "unit": "Mrays/s", "h2d_bytes_per_step": 144, "d2h_bytes_per_step": rays * 3},
            "gpu_launches": int(launches),
            "clocks": sampler.result(),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": rays * bytes_per_ray,
                         "peak_source": peak_src, "algorithmic_bytes_per_ray": bytes_per_ray, "kernel": "trace_kernel<false,false> (STATS off, no-mesh specialisation)"},
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import orc  # checker / baseline only
            S = orc.Scene(rec, 0.01)
            y0 = (H - SAMPLE_ROWS) // 2
            r = S.render(orc.params_from_grt(P), K=7, rect=(0, y0, W, SAMPLE_ROWS), want_float=False, want_flags=False)
            band = r["rgb8"][y0:y0 + SAMPLE_ROWS]
            gpu_band = frame[y0:y0 + SAMPLE_ROWS].cpu().numpy()
            line["cpu_baseline"] = {
                "value": W * SAMPLE_ROWS / r["stats"]["seconds"] / 1e6, "unit": "Mrays/s", "cores": int(r["stats"]["threads"]), "kind": "port",
                "sample": f"{SAMPLE_ROWS} centred full-width rows ({W * SAMPLE_ROWS} rays) of the same frame, K=7",
                "max_lsb_vs_gpu": int(np.abs(band.astype(int) - gpu_band.astype(int)).max()),
                "hits_gathered_per_ray": r["stats"]["hits_gathered"] / r["stats"]["rays"],
            }
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
