"""Where the end-to-end step (host scene in, host frame out) spends its time, per API call.

    PYTHONPATH=. python tools/e2e_breakdown.py [--workload C2] [--spp 16] [--reps 3]
"""
import argparse
import time

import numpy as np
import torch

from gpu_pathtracer_b200.api import PathTracer
from gpu_pathtracer_b200.scene import make_scene


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C2")
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--pageable", action="store_true", help="ordinary numpy arrays, not page-locked")
    args = ap.parse_args()
    sc = make_scene(args.workload)
    sc.params["pathtracing_sample_size"] = args.spp
    pt = PathTracer(device=0)
    cam = sc.camera
    out = None
    if not args.pageable:
        sc = sc.pinned()
        out = torch.empty((cam["height"], cam["width"], 3), dtype=torch.float32, pin_memory=True).numpy()
    rows = []
    for rep in range(args.reps + 1):
        t = [time.perf_counter()]
        pt.set_scene(sc); torch.cuda.synchronize(); t.append(time.perf_counter())
        pt.build(); torch.cuda.synchronize(); t.append(time.perf_counter())
        img = pt.render_pathtrace(out=out); t.append(time.perf_counter())
        st = pt.stats()
        if rep:
            rows.append([(b - a) * 1e3 for a, b in zip(t, t[1:])] + [st["render_ms"]])
    r = np.array(rows).mean(0)
    print(f"{args.workload} spp {args.spp}: set_scene {r[0]:.2f} ms, build {r[1]:.2f} ms "
          f"(build_all_ms {pt.stats()['build_all_ms']:.2f}), render call {r[2]:.2f} ms "
          f"(device frame {r[3]:.2f} ms), total {r[:3].sum():.2f} ms")


if __name__ == "__main__":
    main()
