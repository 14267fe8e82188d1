"""Where does a 1/N-frame shard spend its time?  One GPU renders rank 0's tiles of an N-way
split of the C2 frame with 1, 2, 4, 8 wavefront groups, then once with stage timing (G = 1).
Usage: PYTHONPATH=. python tools/shard_probe.py [--nranks 8] [--spp 16] [--groups 1,2,4,8]
       [--frames 1]  (--frames 1 --groups 4 --no-stage is the form to run under ncu)"""
import argparse
import hashlib
import json
import time

import torch

from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

ap = argparse.ArgumentParser()
ap.add_argument("--nranks", type=int, default=8)
ap.add_argument("--spp", type=int, default=16)
ap.add_argument("--groups", default="1,2,4,8")
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--no-stage", action="store_true")
ap.add_argument("--workload", default="C2")
ap.add_argument("--pipeline", default="rounds,stream")
a = ap.parse_args()
sc = S.make_scene(a.workload, spp=a.spp)
pt = PathTracer(sc)
pt.build()
fb = torch.zeros((sc.camera["height"], sc.camera["width"], 3), dtype=torch.float32, device="cuda")


def run(nr, label):
    for _ in range(a.warm):
        pt.render_pathtrace_device(fb.data_ptr(), 0, nr)
    pt.synchronize()
    t0 = time.perf_counter()
    for _ in range(a.frames):
        pt.render_pathtrace_device(fb.data_ptr(), 0, nr)
    pt.synchronize()
    ms = (time.perf_counter() - t0) / a.frames * 1e3
    st = pt.stats()
    rec = {"nranks": nr, **label, "frame_ms": round(ms, 2), "rays": st["rays_traced"],
           "Mrays_s": round(st["rays_traced"] / ms / 1e3, 1), "launches": st["kernel_launches"],
           "frame_sha1": hashlib.sha1(fb.cpu().numpy().tobytes()).hexdigest()[:12]}
    return rec, st


for nr in sorted({1, a.nranks}):
    for pipe in a.pipeline.split(","):
        pt.set_pipeline(pipe)
        for g in [int(x) for x in a.groups.split(",")] if pipe == "rounds" else [0]:
            pt.set_profiling(time_stages=False)
            pt.set_wavefront_groups(g)
            rec, st = run(nr, {"pipeline": pipe, "groups": g})
            rec["trace_launches"] = st["trace_launches"]
            print(json.dumps(rec), flush=True)
    pt.set_pipeline("rounds")
    if not a.no_stage:
        pt.set_profiling(time_stages=True)
        rec, st = run(nr, {"groups": "stage-timed (1)"})
        rec["trace_ms"] = round(st["trace_kernel_ms"], 2)
        rec["shade_sort_ms"] = round(st["shade_kernel_ms"], 2)
        rec["trace_launches"] = st["trace_launches"]
        print(json.dumps(rec), flush=True)
