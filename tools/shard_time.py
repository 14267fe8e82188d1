"""Time one rank's share of the C2 frame on ONE GPU for every rank of an N-way tile split:
shows load imbalance between ranks and the per-frame fixed cost that bounds strong scaling.
Usage: PYTHONPATH=. python tools/shard_time.py [--nranks 4] [--spp 16]"""
import argparse
import json
import time

import torch

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

ap = argparse.ArgumentParser()
ap.add_argument("--nranks", type=int, default=4)
ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--workload", default="C2")
ap.add_argument("--only-rank0", action="store_true")
ap.add_argument("--rng-mode", type=int, default=0)
ap.add_argument("--lanes", type=int, default=0)
a = ap.parse_args()
sc = S.make_scene(a.workload, spp=a.spp)
sc.params["rng_mode"] = a.rng_mode
pt = PathTracer(sc)
pt.build()
pt.set_fast_rng_lanes(a.lanes)
fb = torch.zeros((sc.camera["height"], sc.camera["width"], 3), dtype=torch.float32, device="cuda")
out = []
for nr in sorted({1, a.nranks}):
    for r in range(1 if a.only_rank0 else nr):
        for timing in (False, True):
            pt.set_profiling(time_stages=timing)
            for _ in range(2):
                pt.render_pathtrace_device(fb.data_ptr(), r, nr)
            pt.synchronize()
            t0 = time.perf_counter()
            for _ in range(3):
                pt.render_pathtrace_device(fb.data_ptr(), r, nr)
            pt.synchronize()
            ms = (time.perf_counter() - t0) / 3 * 1e3
            st = pt.stats()
            rec = {"nranks": nr, "rank": r, "stage_timing": timing, "frame_ms": round(ms, 2),
                   "rays": st["rays_traced"]}
            if timing:
                rec["trace_ms"] = round(st["trace_kernel_ms"], 2)
                rec["shade_ms"] = round(st["shade_kernel_ms"], 2)
                rec["launches"] = st["kernel_launches"]
            out.append(rec)
            print(json.dumps(rec), flush=True)
