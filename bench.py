#!/usr/bin/env python
"""bench.py -- headline measurement of the hot path (contract: see DESIGN.md "Measurement").

A "step" is one pass of the hot path over one batch of synthetic input on the workload
BASELINE.json quotes the metric on (config[1]: 1M-triangle mixed-material scene, 1080p).
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

L2_FLUSH_BYTES = 512 << 20


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "C3", "C4", "C5", "small"])
    ap.add_argument("--photons", type=int, default=0, help="C4: photons per pass override")
    ap.add_argument("--mode", default="auto", choices=["auto", "trace", "render"])
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel")
    ap.add_argument("--pipeline", default="default", choices=["default", "rounds", "stream"],
                    help="gpt_set_pipeline: rounds of launches / two persistent kernels")
    ap.add_argument("--cpu-budget-s", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true",
                    help="skip the run of the reference's own CUDA binary after the timed region")
    ap.add_argument("--no-fast-rng", action="store_true", help="skip the config.fast_rng leg")
    ap.add_argument("--no-also", action="store_true", help="skip the short C4 (SPPM) run appended as 'also'")
    ap.add_argument("--fast-rng", type=int, default=0,
                    help="rng_mode 1 with this many sample lanes per pixel (0 = reference streams)")
    return ap.parse_args()


def make_workload(name: str, spp: int = 0):
    from gpu_pathtracer_b200 import scene as S
    if name == "C2":
        sc = S.make_scene("C2")
    elif name == "C1":
        sc = S.make_scene("C1")
    elif name == "C3":
        sc = S.make_scene("C3")
    elif name == "C5":  # 10 M triangles + fog + SSS at 4K (BASELINE config 5; pass --spp to shorten)
        sc = S.make_scene("C5")
    else:
        sc = S.make_scene("small", n=64, width=256, height=256)
    if spp:
        sc.params["pathtracing_sample_size"] = spp
    return sc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.path = f"/tmp/bench_clocks_{os.getpid()}.csv"
        self.p = None
        try:
            self.f = open(self.path, "w")
            self.p = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self) -> dict:
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.close()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for n, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        try:
            os.remove(self.path)
        except OSError:
            pass
        if sm:
            top = sorted(sm)[len(sm) // 2:]  # samples under load = upper half
            out["sm_mhz"] = float(np.median(top))
            out["sm_max_mhz"] = float(max(mx))
        out["reasons"] = sorted(reasons)
        out["samples"] = len(sm)
        return out


def workload_text(sc, args):
    p, c = sc.params, sc.camera
    return (f"{args.workload}: {sc.num_triangles} triangles, {sc.materials['num_materials']} "
            f"materials, {c['width']}x{c['height']} @ {p['pathtracing_sample_size']} spp x "
            f"{p['dof_sample_size']} dof, level {p['pathtracing_level']}, path tracing")


def reference_window(sc, budget_s: float, nthreads: int):
    """Times the reference's own path_tracing_render (host replay of its headers,
    oracle/_ref) on a bounded window of the SAME frame at the SAME spp, growing the window
    until about `budget_s` seconds of CPU work were spent.  Returns Mrays/s."""
    from tests.refreplay import RefReplay, available
    if not available():
        return None
    ref = RefReplay(sc)
    H, W = sc.camera["height"], sc.camera["width"]
    rays, ms, rows_done = 0, 0.0, 0
    windows = []  # (r0, r1, c0, c1, reference pixels): kept for the parity block
    # horizontal strips around the image centre (terrain + spheres), alternating below / above;
    # a strip has one row per host thread, because the replay hands out work by rows
    strip = max(1, nthreads)
    cols = (W // 4, 3 * W // 4)
    k = 0
    while ms < budget_s * 1e3:
        off = (k // 2) * strip
        r0 = H // 2 + off if k % 2 == 0 else H // 2 - off - strip
        if r0 < 0 or r0 + strip > H:
            break
        px = ref.render_pathtrace(window=(r0, r0 + strip, cols[0], cols[1]), nthreads=nthreads)
        windows.append((r0, r0 + strip, cols[0], cols[1], px[r0:r0 + strip, cols[0]:cols[1]].copy()))
        rays += ref.render_rays
        ms += ref.render_ms
        rows_done += strip
        k += 1
    npx = rows_done * (cols[1] - cols[0])
    return dict(mrays=rays / ms / 1e3, rays=rays, ms=ms, windows=windows,
                frame_fraction=npx / float(H * W),
                sample=f"{rows_done} image rows x columns [{cols[0]},{cols[1]}) around the centre "
                       f"of the same frame at full spp = {100.0 * npx / (H * W):.2f} % of its pixels "
                       f"({rays} rays, {ms / 1e3:.1f} s)")


# Stated tolerance of the bench's own parity check: the GPU frame of this run
# against the reference replay's pixels of the same window, identical per-pixel RNG streams.
# A pixel's dof x spp samples share ONE RNG stream: when a last-ulp difference between CUDA's and
# glibc's libm flips one discrete decision (lobe choice, texel, Russian-roulette-like branch) in
# sample k, every later sample of that pixel draws other numbers and the pixel becomes an
# independent Monte-Carlo estimate of the same integral.  At 2-4 spp that is < 0.1 % of the pixels
# (tests/test_gpu_render.py), at C2's 64 spp x ~2.6 vertices it is ~10 % (measured r02b: 90.6 %
# within 1e-3, 80.8 % bit-equal, RMSE 11 % of the mean = the per-pixel MC noise of those pixels,
# mean luminance equal to 4.5e-4; C3's 8 x 8 samples through nested glass: 85.6 % within 1e-3,
# mean equal to 8e-4; C1: 99.0 %, 2e-5).  The bounds below are those measurements with a margin;
# the unbiasedness bound (mean luminance) is the tight one.
PARITY_MIN_FRAC_1E3 = 0.80
PARITY_MAX_RMSE_FRAC = 0.20
PARITY_MAX_REL_LUM = 5e-3


def parity_block(frame, windows):
    """frame: (H, W, 3) host array of the GPU arm; windows: reference_window()['windows']."""
    got = np.concatenate([frame[r0:r1, c0:c1].reshape(-1, 3) for r0, r1, c0, c1, _ in windows])
    want = np.concatenate([px.reshape(-1, 3) for *_, px in windows])
    close = (np.abs(got - want) <= 1e-3 * (1.0 + np.abs(want))).all(axis=-1)
    mean_w = float(want.mean())
    rmse = float(np.sqrt(np.mean((got.astype(np.float64) - want) ** 2)))
    rel_lum = abs(float(got.mean()) - mean_w) / max(abs(mean_w), 1e-30)
    rows = sum(r1 - r0 for r0, r1, *_ in windows)
    c0, c1 = windows[0][2], windows[0][3]
    out = {
        "against": "reference replay (oracle/_ref/libref_replay.so), same per-pixel RNG streams",
        "window": f"{rows} rows x columns [{c0},{c1}) around the image centre, {got.shape[0]} pixels",
        "frac_within_1e-3": float(close.mean()), "frac_bit_equal": float((got == want).all(axis=-1).mean()),
        "rmse": rmse, "rmse_over_mean": rmse / max(abs(mean_w), 1e-30), "rel_mean_lum": rel_lum,
        "finite": bool(np.isfinite(got).all()),
        "tolerance": {"frac_within_1e-3_min": PARITY_MIN_FRAC_1E3,
                      "rmse_over_mean_max": PARITY_MAX_RMSE_FRAC, "rel_mean_lum_max": PARITY_MAX_REL_LUM},
    }
    out["ok"] = bool(out["finite"] and out["frac_within_1e-3"] >= PARITY_MIN_FRAC_1E3
                     and out["rmse_over_mean"] <= PARITY_MAX_RMSE_FRAC and rel_lum <= PARITY_MAX_REL_LUM)
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU code on all host cores (rank 0 only)."""
    if rank != 0:
        return
    if args.workload == "C4":
        print(json.dumps({"impl": "reference", "unavailable":
                          "the reference arm is wired for the headline workload (C2); SPPM is "
                          "compared in tools/run_reference_cuda_sppm.py and tests/test_gpu_sppm.py"}))
        return
    sc = make_workload(args.workload, args.spp)
    cores = os.cpu_count() or 1
    times, rays = [], []
    sample, frac = "", 0.0
    for s in range(args.warmup + args.steps):
        r = reference_window(sc, 6.0 if s >= args.warmup else 1.0, cores)
        if s >= args.warmup:
            times.append(r["ms"]); rays.append(r["rays"]); sample = r["sample"]
            frac = r["frame_fraction"]
    ms_per_step = float(np.mean(times))
    val = float(np.sum(rays) / np.sum(times) / 1e3)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_text(sc, args)},
        "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                         "frame_fraction": frac, "sample": "per step: " + sample},
        "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def ref_cuda_block(sc, workload: str, rays: float, our_ms: float) -> dict:
    """north_star's GPU baseline: the reference's OWN CUDA build (unmodified sources + the jpeg
    stub, the reference Makefile's flags with SM=100; oracle/_ref/gpu_path_tracing_100, built by
    `make -C oracle refcuda`) run on the same scene written to OBJ/MTL/master files, on the same
    GPU, after the timed region.  Its "Rendering" stage (print_end_process around
    path_tracing_render + sync, src/main.cu:1266) against our frame time; the ray count is
    ours, which equals the reference's traverse_bvh(node_list[0]) call count sample for sample."""
    import re
    import shutil
    import tempfile
    from gpu_pathtracer_b200 import scene as S
    binary = os.path.join(ROOT, "oracle", "_ref", "gpu_path_tracing_100")
    if not os.path.exists(binary):
        return {"unavailable": "oracle/_ref/gpu_path_tracing_100 not built (make -C oracle refcuda)"}
    folder = tempfile.mkdtemp(prefix="gpt_refcuda_")
    try:
        master = S.write_scene_files(sc, folder, workload)
        # The reference times a stage with clock() (CPU time, util/string_util.h:90-111), which
        # under-reports whenever cudaDeviceSynchronize blocks instead of spinning.  Its stdout is
        # therefore line-buffered (stdbuf) and the arrival of its own "Started process" / "Done
        # with process" lines of the Rendering stage is timed on the wall clock here.
        cmd = [binary, master, os.path.join(folder, "ref_out")]
        if shutil.which("stdbuf"):
            cmd = ["stdbuf", "-oL"] + cmd
        t0 = time.perf_counter()
        proc = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, bufsize=1)
        t_start = t_done = None
        cpu_clock_s, last_done, tail = None, None, []
        for ln in proc.stdout:
            now = time.perf_counter()
            tail = (tail + [ln.rstrip()])[-8:]
            if ln.startswith("Started process : Rendering"):
                t_start = now
            elif ln.startswith("Done with process : "):
                last_done = ln[len("Done with process : "):].strip()
                if last_done == "Rendering":
                    t_done = now
            elif ln.startswith("The process took  :") and last_done == "Rendering":
                m = re.search(r"([0-9.]+) seconds", ln)
                cpu_clock_s = float(m.group(1)) if m else None
                last_done = None
            if now - t0 > 900:
                proc.kill()
                return {"unavailable": "reference CUDA binary did not finish in 900 s"}
        rc_ref = proc.wait()
        wall = time.perf_counter() - t0
        if rc_ref != 0:
            return {"unavailable": f"reference CUDA binary exited with {rc_ref}: " + " | ".join(tail)[-300:]}
        if t_start is None or t_done is None:
            return {"unavailable": "no 'Rendering' stage lines in the reference binary's output"}
        render_s = t_done - t_start
        return {"binary": "oracle/_ref/gpu_path_tracing_100 (reference sources unmodified + jpeg "
                          "stub, -O3 -gencode arch=compute_100,code=sm_100, default FMA)",
                "render_s": render_s, "render_cpu_clock_s": cpu_clock_s, "wall_s": wall,
                "value": rays / render_s / 1e6, "unit": "Mrays/s",
                "ours_over_ref_cuda": (render_s * 1e3) / our_ms,
                "note": "same scene files, same GPU, run after the timed region; rays = our "
                        "closest-hit count (identical paths); render_s = wall clock between the "
                        "reference's own 'Started process : Rendering' and 'Done with process : "
                        "Rendering' lines (path_tracing_render + cudaDeviceSynchronize, "
                        "src/main.cu:1264-1280; stdout line-buffered); render_cpu_clock_s = the "
                        "clock() seconds it prints itself"}
    finally:
        shutil.rmtree(folder, ignore_errors=True)


def newest_traffic_capture(workload: str):
    """profiles/*_trace_traffic_<workload>.json written by tools/trace_traffic.py from an ncu
    metrics pass over every trace launch of whole frames; None when there is no capture."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", f"*_trace_traffic_{workload}.json")))
    if not files:
        return None, None
    try:
        return json.load(open(files[-1])), os.path.relpath(files[-1], ROOT)
    except Exception:
        return None, None


def newest_trace_capture_counters():
    """Issue-slot utilisation, lanes per instruction and occupancy of the trace kernel from the
    newest `ncu --set full` capture under profiles/ (*_k_pt_trace_raw.csv, first launch): the
    kernel is issue-bound, so this -- not a byte count -- says how close it is to its ceiling."""
    import csv
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_k_pt_trace_raw.csv")))
    if not files:
        return None
    try:
        rows = list(csv.reader(open(files[-1])))
        hdr, row = rows[0], rows[2]
        get = lambda k: float(row[hdr.index(k)].replace(",", ""))
        return {"issue_slots_busy_frac": get("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
                "lanes_per_instruction": get("smsp__thread_inst_executed_per_inst_executed.ratio"),
                "warps_active_frac": get("sm__warps_active.avg.pct_of_peak_sustained_active") / 100.0,
                "kernel": row[hdr.index("Kernel Name")].split("(")[0].split("::")[-1][:40],
                "source": os.path.relpath(files[-1], ROOT)}
    except Exception:
        return None


def scene_bytes(sc) -> int:
    n = 0
    for d in (sc.mesh, sc.materials, sc.bg):
        for v in d.values():
            if isinstance(v, np.ndarray):
                n += v.nbytes
    return n


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gpu_pathtracer_b200.api import PathTracer
    from gpu_pathtracer_b200.parallel import init_comm

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    sc = make_workload(args.workload, args.spp)
    H, W = sc.camera["height"], sc.camera["width"]
    pt = PathTracer(sc, device=local_rank)
    pt.set_pipeline(args.pipeline)
    if args.fast_rng:
        pt.set_params(rng_mode=1)
        pt.set_fast_rng_lanes(args.fast_rng)
    # a real (non-legacy) stream: the library launches on it and the events below time it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    pt.set_stream(stream.cuda_stream)
    init_comm(pt, rank, world)  # libgpt.so's own NCCL communicator (gpt_comm_init)
    build_stats = pt.build()
    build_ms = min(pt.rebuild_main_bvh() for _ in range(5))
    fb = torch.zeros((H, W, 3), dtype=torch.float32, device=dev)
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if world > 1:
        # NCCL sets its connections up lazily at a communicator's first collective (~0.2 s): that is
        # start-up, not part of a frame, so the library's communicator does one untimed reduce
        pt.gather_frame(fb.data_ptr(), 0)
        pt.synchronize()
        fb.zero_()

    # ---- device-resident timing: scene + BVH already in HBM, K full-frame renders of this
    #      rank's tiles, L2 flushed between steps, CUDA events on the launching stream
    for _ in range(args.warmup):
        flush.zero_()
        pt.render_pathtrace_device(fb.data_ptr(), rank, world)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    launches, rays = 0, 0
    for a, b in evs:
        flush.zero_()
        a.record(stream)
        pt.render_pathtrace_device(fb.data_ptr(), rank, world)
        b.record(stream)
        st = pt.stats()
        launches += st["kernel_launches"]
        rays = st["rays_traced"]
    barrier()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(step_ms))
    clocks = sampler.stop() if sampler else None

    # ---- the same frames with rng_mode 1 (8 independent sample lanes per pixel: NOT the
    #      reference's sample order, see config.fast_rng) -- reported beside the headline, which
    #      stays on the reference's streams
    fast_ms, fast_rays, fast_lanes = 0.0, 0, 8
    if not args.fast_rng and not args.no_fast_rng:
        pt.set_params(rng_mode=1)
        pt.set_fast_rng_lanes(fast_lanes)
        try:
            for _ in range(2):
                flush.zero_()
                pt.render_pathtrace_device(fb.data_ptr(), rank, world)
            barrier()
            fevs = []
            for _ in range(min(args.steps, 5)):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(stream)
                pt.render_pathtrace_device(fb.data_ptr(), rank, world)
                b.record(stream)
                fevs.append((a, b))
                fast_rays = pt.stats()["rays_traced"]
            barrier()
            fast_ms = float(np.mean([a.elapsed_time(b) for a, b in fevs]))
        finally:
            pt.set_params(rng_mode=0)  # everything below is on the reference's streams again

    # ---- profiling passes (not timed): per-stage kernel time, then visit counts
    pt.set_profiling(time_stages=True)
    pt.render_pathtrace_device(fb.data_ptr(), rank, world)
    st_t = pt.stats()
    pt.set_profiling(count_visits=True)
    pt.render_pathtrace_device(fb.data_ptr(), rank, world)
    st_c = pt.stats()
    pt.set_profiling()
    v_int, v_leaf = st_c["internal_visits"], st_c["leaf_tests"]
    v_wide = st_c["wide_node_visits"]  # the share of v_int made on 48-byte nodes of the 4-wide tree
    alg_bytes = 48.0 * rays + 64.0 * (v_int - v_wide) + 48.0 * v_wide + 48.0 * v_leaf
    trace_ms, shade_ms = st_t["trace_kernel_ms"], st_t["shade_kernel_ms"]
    n_trace = max(1, st_t["trace_launches"])

    # ---- end to end through the reference-facing C-ABI with HOST buffers: upload the scene
    #      SoA, build, render, read the frame back (NCCL sum = tile gather when N > 1)
    #      The host arrays are page-locked (Scene.pinned, as the bench contract asks), the
    #      frame lands in a page-locked array too.
    e2e_steps = max(1, min(args.steps, 3))
    sc_host = sc.pinned()
    frame_host = torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        pt.set_scene(sc_host)
        pt.build()
        if world > 1:
            fb.zero_()
            pt.render_pathtrace_device(fb.data_ptr(), rank, world)
            pt.gather_frame(fb.data_ptr(), 0)  # gpt_gather_frame: ncclReduce on the context's stream
            if rank == 0:
                frame_host.copy_(fb.view(H, W, 3))
                torch.cuda.current_stream().synchronize()
        else:
            img = pt.render_pathtrace(out=frame_host.numpy())
    barrier()
    e2e_ms = (time.perf_counter() - t0) / e2e_steps * 1e3

    t = torch.tensor([total_ms, e2e_ms, fast_ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(rays), float(fast_rays)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms, fast_ms = t.tolist()
    rays_all, fast_rays_all = tot.tolist()
    if rank != 0:
        return
    # the full frame of the e2e leg on the host (N > 1: after gpt_gather_frame on rank 0)
    frame = frame_host.numpy() if world > 1 else img
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ms_per_step = total_ms / args.steps
    value = rays_all / (ms_per_step * 1e-3) / 1e6
    spp_total = sc.params["pathtracing_sample_size"] * sc.params["dof_sample_size"]
    achieved = alg_bytes / (trace_ms * 1e-3) / 1e9 if trace_ms > 0 else 0.0
    cap, cap_file = newest_traffic_capture(args.workload)
    rays_per_launch = rays / max(1, n_trace)
    roofline = {
        "bound": "hbm",
        "kernel": "k_pt_trace_wide" if 2 * v_wide >= v_int else "k_pt_trace",
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "what": "ALGORITHMIC bytes (SURVEY 8d: 48 B per ray + 48 B per visit of a 4-wide node + "
                "64 B per visit of a binary LBVH node + 48 B per triangle test, from the kernels' "
                "own visit counters -- the binary walk's are pinned against a host count in "
                "tests/test_gpu_counters.py) / summed event time of the trace launches of a "
                "stage-timed frame.  On C1-C4 the tree is L2 resident, so this is an L2-served "
                "proxy, not HBM headroom: physical_* are the DRAM bytes ncu counted",
        "wide_share_of_node_visits": v_wide / max(1, v_int),
        "launches": int(n_trace), "algorithmic_bytes_per_launch": alg_bytes / n_trace,
        "avg_launch_ms": trace_ms / n_trace,
        "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback",
        # physical bytes: ncu metrics pass over EVERY trace launch of whole frames of this
        # workload (tools/trace_traffic.py), per ray, scaled to this run's rays per launch
        "traffic": (cap["dram_bytes_per_ray"] * rays_per_launch) if cap else None,
        "traffic_source": cap_file,
    }
    ncu_counters = newest_trace_capture_counters()
    if ncu_counters:
        roofline["ncu"] = ncu_counters
    if cap:
        roofline["physical_dram_gbs"] = cap["dram_bytes_per_ray"] * rays / (trace_ms * 1e-3) / 1e9
        roofline["physical_frac"] = roofline["physical_dram_gbs"] / peak
        roofline["l2_gbs"] = cap["l2_bytes_per_ray"] * rays / (trace_ms * 1e-3) / 1e9
        roofline["l2_bytes_per_ray"] = cap["l2_bytes_per_ray"]
        roofline["dram_bytes_per_ray"] = cap["dram_bytes_per_ray"]
    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {
            "workload": workload_text(sc, args),
            "step": "one full frame: ray-gen + trace + shade wavefront over all pixels/samples",
            "rays": "closest-hit traversals of the main BVH (primary + bounce), as the "
                    "reference's traverse_bvh(node_list[0]) call count",
            "cache": "L2 flushed (512 MiB write) between timed steps",
            "sharding": "scene replicated, 32x32 pixel tiles dealt round-robin to ranks",
            "pipeline": {"default": "rounds (library default)"}.get(args.pipeline, args.pipeline),
            "rng": ("reference streams (seed 1984, seq = pixel; rng_mode 0)" if not args.fast_rng else
                    f"rng_mode 1: {args.fast_rng} independent sample lanes per pixel (NOT the "
                    "reference's sample order; statistical parity only)"),
            "fast_rng": ({"what": f"rng_mode 1, {fast_lanes} independent sample lanes per pixel "
                                  "(seq = pixel * L + lane): same estimator, other random numbers "
                                  "than the reference -- statistical parity only (tests/test_gpu_"
                                  "render.py::test_fast_rng_mode); max over ranks, L2 flushed",
                          "value": fast_rays_all / (fast_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                          "ms_per_step": fast_ms} if fast_ms > 0 else None),
            "spp_per_s": spp_total / (ms_per_step * 1e-3),
            "rays_per_frame": rays_all,
            "lbvh_build_ms": build_ms, "build_all_ms": build_stats["build_all_ms"],
            # 288 algorithmic bytes per triangle against the measured HBM peak
            "lbvh_build_roofline_frac": 288.0 * sc.num_triangles / (build_ms * 1e-3) / 1e9 / peak,
            "rank0_stage_ms": {"trace": trace_ms, "shade": shade_ms,
                               "frame": float(np.mean(step_ms))},
            "avg_internal_visits_per_ray": v_int / max(1, rays),
            "avg_leaf_tests_per_ray": v_leaf / max(1, rays),
        },
        "gpu_launches": int(launches),
        "e2e": {"value": rays_all / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "h2d_bytes_per_step": scene_bytes(sc), "d2h_bytes_per_step": H * W * 12,
                "what": "gpt_set_mesh/materials/textures + gpt_build + gpt_render_pathtrace "
                        "with page-locked host arrays in, page-locked host frame out"},
        "roofline": roofline,
        "clocks": clocks,
    }
    rc = 0
    if not args.no_cpu_baseline:
        # the reference's own path_tracing_render on a window of the SAME frame: timed on one host
        # thread at N = 1 (cpu_baseline), on all host threads otherwise; its pixels are kept and
        # compared with the same window of the GPU frame of this run (parity)
        nthreads = 1 if world == 1 else (os.cpu_count() or 1)
        cb = reference_window(sc, args.cpu_budget_s, nthreads)
        if cb:
            if world == 1:
                line["cpu_baseline"] = {
                    "value": cb["mrays"], "unit": "Mrays/s", "cores": 1, "kind": "reference",
                    "frame_fraction": cb["frame_fraction"],
                    "sample": cb["sample"] + f", 1 thread of {os.cpu_count()} host cores"}
            if not args.fast_rng:
                line["parity"] = parity_block(frame, cb["windows"])
                if not line["parity"]["ok"]:
                    rc = 3
    if world == 1 and not args.no_ref_cuda and args.workload in ("C1", "C2", "C3"):
        try:
            line["ref_cuda"] = ref_cuda_block(sc, args.workload, rays_all, ms_per_step)
        except Exception as exc:  # noqa: BLE001 -- a reported baseline, never worth the line
            line["ref_cuda"] = {"unavailable": f"{type(exc).__name__}: {exc}"[:300]}
    line["_rc"] = rc
    return line


def run_sppm(args, rank, world, local_rank):
    """--workload C4: stochastic progressive photon mapping, 10 M photons per pass, 1080p.  A
    step = one SPPM pass (photon trace of this rank's photon range, photon trees, eye pass,
    gather, the NCCL all-reduce of the per-hit-point accumulators when N > 1, update)."""
    import torch
    import torch.distributed as dist
    from gpu_pathtracer_b200 import scene as S
    from gpu_pathtracer_b200.api import PathTracer
    from gpu_pathtracer_b200.parallel import init_comm, photon_range

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    photons = args.photons or 10_000_000
    sc = S.make_scene("C4", photons=photons, passes=args.warmup + args.steps)
    pt = PathTracer(sc, device=local_rank)
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    pt.set_stream(stream.cuda_stream)
    init_comm(pt, rank, world)  # libgpt.so's own NCCL communicator (gpt_comm_init)
    pt.build()
    pt.sppm_init()
    b, e = photon_range(photons, rank, world)
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def one_pass(p):
        pt.sppm_pass(p, b, e)
        r = pt.stats()["rays_traced"]
        pt.sppm_allreduce()  # gpt_sppm_allreduce: stream-ordered, a no-op with one rank
        pt.sppm_update(p)
        return r

    for p in range(args.warmup):
        flush.zero_()
        one_pass(p)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    evs, rays, launches = [], 0, 0
    for k in range(args.steps):
        flush.zero_()
        a_, b_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a_.record(stream)
        rays += one_pass(args.warmup + k)
        b_.record(stream)
        launches += pt.stats()["kernel_launches"] + 1
        evs.append((a_, b_))
    barrier()
    total_ms = float(sum(x.elapsed_time(y) for x, y in evs))
    clocks = sampler.stop() if sampler else None
    # stage times + visit counts of one more pass (not timed)
    p_extra = args.warmup + args.steps
    stage = {}
    pt.set_profiling(count_visits=True)
    v_int = v_leaf = prof_rays = eye_rays = 0
    for name, mask in (("photon", 1), ("eye", 2), ("gather", 4)):
        pt.sppm_pass_stages(p_extra, mask, b, e)
        st = pt.stats()
        stage[name] = st["render_ms"]
        if name == "photon":
            v_int, v_leaf, prof_rays = st["internal_visits"], st["leaf_tests"], st["rays_traced"]
        if name == "eye":
            eye_rays = st["rays_traced"]
    pt.set_profiling()
    pt.sppm_pass_stages(p_extra, 1, b, e)
    photon_ms = pt.stats()["render_ms"]
    stats = pt.sppm_stats()
    # e2e: host scene in, host frame out, for `steps` passes
    sc_host = sc.pinned()
    barrier()
    t0 = time.perf_counter()
    pt.set_scene(sc_host)
    pt.build()
    pt.sppm_init()
    e2e_rays = init_rays = pt.stats()["rays_traced"]
    for p in range(args.steps):
        e2e_rays += one_pass(p)
    img = pt.sppm_resolve(0, args.steps)
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3
    t = torch.tensor([total_ms, e2e_ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(rays), float(e2e_rays)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms = t.tolist()
    rays_all, e2e_rays_all = tot.tolist()
    # the eye passes (init included) are sharded by tiles when N > 1 (each rank counts its own
    # pixels' rays); only where they are replicated (GPT_SPPM_SHARD_EYE=0) are their rays counted
    # ONCE, not once per rank
    eye_sharded = world > 1 and os.environ.get("GPT_SPPM_SHARD_EYE", "1") != "0"
    if not eye_sharded:
        rays_all -= (world - 1) * eye_rays * args.steps
        e2e_rays_all -= (world - 1) * (init_rays + eye_rays * args.steps)
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ms_per_step = total_ms / args.steps
    alg = 48.0 * prof_rays + 64.0 * v_int + 48.0 * v_leaf + 52.0 * (stats["n_surface"] + stats["n_volume"])
    achieved = alg / (photon_ms * 1e-3) / 1e9 if photon_ms > 0 else 0.0
    H, W = sc.camera["height"], sc.camera["width"]
    line = {
        "metric": "Mrays/s", "value": rays_all / (total_ms * 1e-3) / 1e6, "unit": "Mrays/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {
            "workload": f"C4: SPPM, {sc.num_triangles} triangles (room + glass sphere + fog box), "
                        f"{W}x{H}, {photons} photons per pass",
            "step": "one SPPM pass: photon trace + photon LBVHs + eye pass + gather (+ NCCL "
                    "all-reduce of 8 floats per pixel when N > 1) + update",
            "rays": "closest-hit traversals of the main BVH: photon, eye, direct-light and "
                    "visibility rays (the replicated eye pass counted once)",
            "cache": "L2 flushed (512 MiB write) between timed steps",
            "sharding": ("scene replicated, photons split by index range; volume photons all-gathered, "
                         "eye pass sharded by 32x32 tiles, hit points all-reduced (12 floats per pixel)"
                         if eye_sharded else "scene + eye pass replicated, photons split by index range"),
            "rank0_stage_ms": stage, "photons_stored": [stats["n_surface"], stats["n_volume"]],
            "mean_radius": stats["mean_radius"],
        },
        "gpu_launches": int(launches),
        "e2e": {"value": e2e_rays_all / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "h2d_bytes_per_step": scene_bytes(sc) // max(1, args.steps),
                "d2h_bytes_per_step": H * W * 12 // max(1, args.steps),
                "what": "gpt_set_* + gpt_build + gpt_sppm_init + passes + gpt_sppm_resolve, host "
                        "arrays in, host frame out"},
        "roofline": {"bound": "hbm", "kernel": "photon stage: k_photon_step + k_ph_trace", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                     "algorithmic_bytes_per_launch": alg, "avg_launch_ms": photon_ms,
                     "note": "photon stage = wavefront photon trace (k_photon_step, k_ph_trace, coherence sort) "
                             "+ compaction + two photon LBVH builds",
                     "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
        "clocks": clocks,
    }
    return line


def also_block(args, rank, world, local_rank):
    """A short run of config 4 (SPPM, 10 M photons per pass, the per-pass NCCL all-reduce and the
    sharded eye pass exercised when N > 1) appended to the headline line as "also", so that the
    driver's 1 / 2 / 4 / 8-GPU runs carry a measured SPPM number too."""
    import copy
    a = copy.copy(args)
    a.steps, a.warmup, a.workload = 4, 2, "C4"
    line = run_sppm(a, rank, world, local_rank)
    if line is None:
        return None
    return {"C4": {"workload": line["config"]["workload"], "pass_ms": line["ms_per_step"],
                   "Mrays/s": line["value"], "e2e_Mrays/s": line["e2e"]["value"], "nranks": world,
                   "steps": a.steps, "warmup": a.warmup, "sharding": line["config"]["sharding"],
                   "rank0_stage_ms": line["config"]["rank0_stage_ms"],
                   "photon_stage_roofline_frac": line["roofline"]["frac"]}}


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local_rank)
        # NCCL prints its version banner on STDOUT when the communicator comes up (NCCL_DEBUG =
        # VERSION on the pool's boxes); the contract is ONE JSON line there, so stdout points at
        # stderr until the first collective has run
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", rank=rank, world_size=world,
                                    device_id=torch.device(f"cuda:{local_rank}"))
            warm = torch.zeros(1, device=f"cuda:{local_rank}")
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    rc = 0
    try:
        if args.workload == "C4":
            line = run_sppm(args, rank, world, local_rank)
        else:
            line = run_ours(args, rank, world, local_rank)
            if args.workload == "C2" and not args.no_also:
                try:  # an extra: it must never cost the headline line
                    also = also_block(args, rank, world, local_rank)  # every rank takes part
                except Exception as exc:  # noqa: BLE001
                    also = {"error": f"{type(exc).__name__}: {exc}"[:300]}
                if line is not None:
                    line["also"] = also
        if line is not None:
            rc = line.pop("_rc", 0)
            print(json.dumps(line), flush=True)
            if rc:
                sys.stderr.write("bench.py: the GPU frame is outside the stated tolerance of the "
                                 "reference replay's pixels (see \"parity\" in the line)\n")
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()
    if rc:
        raise SystemExit(rc)


if __name__ == "__main__":
    main()
