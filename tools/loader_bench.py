"""Scene packer timing: gpt_scene_files_parse (all host threads, and 1 thread) against the
reference's own loaders (oracle/_ref/libref_loader.so, single-threaded by construction) on the
same files.  Host only.  Usage: PYTHONPATH=. python tools/loader_bench.py [--workload C2]"""
import argparse
import ctypes as C
import json
import os
import tempfile
import time

from gpu_pathtracer_b200 import cabi, scene as S
from gpu_pathtracer_b200.api import load_library

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="C2")
ap.add_argument("--out", default=None)
a = ap.parse_args()
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sc = S.make_scene(a.workload, width=64, height=64)
with tempfile.TemporaryDirectory() as d:
    t0 = time.perf_counter()
    master = S.write_scene_files(sc, d, "w")
    write_s = time.perf_counter() - t0
    lib = load_library()
    lib.gpt_scene_files_free.restype = None
    res = {"workload": a.workload, "triangles": sc.num_triangles, "host_cores": os.cpu_count(),
           "write_files_s": round(write_s, 2)}
    for threads in (0, 1):
        best = None
        for _ in range(3):
            h, v = cabi.parse_scene_files(lib, "gpt_", master, threads)
            rec = (v.parse_ms, v.obj_parse_ms, v.threads, v.obj_bytes)
            lib.gpt_scene_files_free(h)
            best = rec if best is None or rec[0] < best[0] else best
        key = "ours_all_threads" if threads == 0 else "ours_1_thread"
        res[key] = {"parse_ms": round(best[0], 1), "obj_ms": round(best[1], 1), "threads": best[2],
                    "obj_MB": round(best[3] / 1e6, 1),
                    "obj_MB_per_s": round(best[3] / 1e6 / (best[1] / 1e3), 1)}
    ref_so = os.path.join(ROOT, "oracle", "_ref", "libref_loader.so")
    if os.path.exists(ref_so):
        ref = C.CDLL(ref_so)
        ref.ref_scene_files_free.restype = None
        devnull = os.open(os.devnull, os.O_WRONLY)
        saved = os.dup(1)
        os.dup2(devnull, 1)  # the reference prints a line per material / usemtl
        try:
            h, v = cabi.parse_scene_files(ref, "ref_", master)
        finally:
            os.dup2(saved, 1)
        res["reference_loader"] = {"parse_ms": round(v.parse_ms, 1), "obj_ms": round(v.obj_parse_ms, 1),
                                   "threads": 1}
        ref.ref_scene_files_free(h)
        res["speedup_all_threads"] = round(v.parse_ms / res["ours_all_threads"]["parse_ms"], 1)
        res["speedup_1_thread"] = round(v.parse_ms / res["ours_1_thread"]["parse_ms"], 1)
print(json.dumps(res))
if a.out:
    open(a.out, "w").write(json.dumps(res, indent=1) + "\n")
