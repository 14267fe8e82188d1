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
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C2", choices=["C1", "C2", "small"])
    ap.add_argument("--mode", default="auto", choices=["auto", "trace", "render"])
    ap.add_argument("--spp", type=int, default=0, help="override samples per pixel")
    ap.add_argument("--cpu-sample-rays", type=int, default=4000000)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def make_workload(name: str, spp: int = 0):
    from gpu_pathtracer_b200 import scene as S
    if name == "C2":
        sc = S.make_scene("C2")
    elif name == "C1":
        sc = S.make_scene("C1")
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


def build_ray_batches(sc, pt, rank: int, world: int):
    """R1 (primary, pixel centres) + R2 (cosine-hemisphere bounce rays from the R1 hit
    points, PCG64 seed 7).  Rays are dealt to ranks in interleaved 4096-ray chunks."""
    from gpu_pathtracer_b200 import rays as R
    r1 = R.camera_rays(sc)
    h = pt.trace(r1)
    hit = h["tri_id"] >= 0
    d = r1[:, 4:7] / np.linalg.norm(r1[:, 4:7], axis=1, keepdims=True)
    pts = r1[hit, 0:3] + h["t"][hit, None] * d[hit]
    nrm = pt.export_triangles()["normal"][h["tri_id"][hit]]
    flip = (nrm * d[hit]).sum(1) > 0
    nrm = np.where(flip[:, None], -nrm, nrm)
    r2 = R.bounce_rays(pts.astype(np.float64), nrm.astype(np.float64), seed=7)
    allr = np.concatenate([r1, r2]).astype(np.float32)
    if world > 1:
        chunk = np.arange(allr.shape[0]) // 4096
        allr = allr[chunk % world == rank]
    return np.ascontiguousarray(allr), int(r1.shape[0]), int(r2.shape[0])


def cpu_baseline_trace(sc, sample_rays: np.ndarray, nthreads: int, budget_s: float):
    """Times the reference's own traverse_bvh + Triangle::hit (oracle/_ref host replay)."""
    from tests.refreplay import RefReplay, available
    if not available():
        return None
    t0 = time.time()
    ref = RefReplay(sc)
    build_s = time.time() - t0
    done, ms = 0, 0.0
    chunk = 4000 * max(1, nthreads)
    while done < sample_rays.shape[0] and ms < budget_s * 1e3:
        part = sample_rays[done:done + chunk]
        ms += ref.trace(part, nthreads=nthreads)["ms"]
        done += part.shape[0]
    return dict(mrays=done / ms / 1e3, rays=done, ms=ms, build_s=build_s)


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU code on the host cores (rank 0 only)."""
    if rank != 0:
        return
    from gpu_pathtracer_b200 import rays as R
    sc = make_workload(args.workload, args.spp)
    cores = os.cpu_count() or 1
    from tests.refreplay import RefReplay
    ref = RefReplay(sc)
    r1 = R.camera_rays(sc)
    rng = np.random.Generator(np.random.PCG64(11))
    per_step = 20000 * cores
    times = []
    for s in range(args.warmup + args.steps):
        idx = rng.integers(0, r1.shape[0], size=per_step)
        ms = ref.trace(r1[idx], nthreads=cores)["ms"]
        if s >= args.warmup:
            times.append(ms)
    ms_per_step = float(np.mean(times))
    val = per_step / ms_per_step / 1e3
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"{args.workload}: {sc.num_triangles} triangles, "
                               f"{sc.camera['width']}x{sc.camera['height']} primary rays",
                   "stage": "trace (traverse_bvh + Triangle::hit)"},
        "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                         "sample": f"{per_step} primary rays per step sampled (PCG64 seed 11) "
                                   f"from the {r1.shape[0]} camera rays, all host threads"},
        "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from gpu_pathtracer_b200.api import PathTracer, GptError

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    sc = make_workload(args.workload, args.spp)
    pt = PathTracer(sc, device=local_rank)
    # a real (non-legacy) stream: the library launches on it and the events below time it
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    pt.set_stream(stream.cuda_stream)
    build_stats = pt.build()
    build_ms = min(pt.rebuild_main_bvh() for _ in range(5))

    rays_np, n_r1, n_r2 = build_ray_batches(sc, pt, rank, world)
    n = rays_np.shape[0]
    d_rays = torch.from_numpy(rays_np).to(dev)
    d_hits = torch.empty((n, 8), dtype=torch.float32, device=dev)
    flush = torch.empty(L2_FLUSH_BYTES // 4, dtype=torch.float32, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing: K steps, L2 flushed between steps, CUDA events on the
    #      stream the kernels are launched on
    for _ in range(args.warmup):
        flush.zero_()
        pt.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr())
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(args.steps)]
    launches = 0
    for a, b in evs:
        flush.zero_()
        a.record(stream)
        pt.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr())
        b.record(stream)
        launches += pt.stats()["kernel_launches"]
    barrier()
    step_ms = [a.elapsed_time(b) for a, b in evs]
    total_ms = float(sum(step_ms))
    clocks = sampler.stop() if sampler else None

    # ---- visit counts for the algorithmic-byte roofline (same traversal, not timed)
    cnt = pt.trace_count_device(d_rays.data_ptr(), n)
    v_int, v_leaf = cnt["internal_visits"], cnt["leaf_tests"]
    alg_bytes = 48.0 * n + 64.0 * v_int + 48.0 * v_leaf

    # ---- end to end through the host C-ABI call: pinned host rays in, host hits out
    e2e_steps = max(1, min(args.steps, 5))
    pt.trace(rays_np[:4096])
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        out = pt.trace(rays_np)
    barrier()
    e2e_s = (time.perf_counter() - t0) / e2e_steps

    t = torch.tensor([total_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(n), alg_bytes, float(v_int), float(v_leaf)], dtype=torch.float64,
                       device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms = t.tolist()
    n_all, bytes_all, vint_all, vleaf_all = tot.tolist()
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    ms_per_step = total_ms / args.steps
    value = n_all / (ms_per_step * 1e-3) / 1e6
    achieved = (alg_bytes / (float(np.mean(step_ms)) * 1e-3)) / 1e9  # this rank's kernel
    line = {
        "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
        "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {
            "workload": f"{args.workload}: {sc.num_triangles} triangles, "
                        f"{sc.camera['width']}x{sc.camera['height']}: {n_r1} primary + {n_r2} "
                        "cosine-bounce rays per step",
            "stage": "trace (closest hit, main BVH)",
            "cache": "L2 flushed (512 MiB write) between timed steps",
            "sharding": "rays interleaved in 4096-ray chunks across ranks, scene replicated",
            "lbvh_build_ms": build_ms, "build_all_ms": build_stats["build_all_ms"],
            "avg_internal_visits_per_ray": vint_all / n_all,
            "avg_leaf_tests_per_ray": vleaf_all / n_all,
        },
        "gpu_launches": int(launches),
        "e2e": {"value": n_all / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s",
                "h2d_bytes_per_step": int(32 * n_all), "d2h_bytes_per_step": int(32 * n_all)},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": None,
                     "kernel": "k_trace", "algorithmic_bytes_per_launch": alg_bytes,
                     "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        "clocks": clocks,
    }
    if not args.no_cpu_baseline:
        rng = np.random.Generator(np.random.PCG64(11))
        sample = rays_np[rng.integers(0, n, size=args.cpu_sample_rays)]
        cb = cpu_baseline_trace(sc, sample, 1, 15.0)
        if cb:
            line["cpu_baseline"] = {
                "value": cb["mrays"], "unit": "Mrays/s", "cores": 1, "kind": "reference",
                "sample": f"{cb['rays']} rays sampled (PCG64 seed 11) from this step's ray "
                          f"batch, single thread of {os.cpu_count()} host cores, "
                          f"{cb['ms'] / 1e3:.1f} s"}
    print(json.dumps(line), flush=True)


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
        dist.init_process_group("nccl", rank=rank, world_size=world)
    try:
        run_ours(args, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
