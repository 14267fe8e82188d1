"""Runs the reference's own CUDA build (oracle/_ref/gpu_path_tracing_100, sm_100) in render_mode 2
on the config-4 scene written to files, next to our SPPM passes on the same scene.

    PYTHONPATH=. python tools/run_reference_cuda_sppm.py [photons] [passes] [cells] [W] [H] [json-out]

Reports per-pass stage seconds of the reference (its own print_end_process lines: photon pass,
photon gather/sort/tree stages, ray tracing pass, update hit point parameters) and our stage
milliseconds, plus image agreement of the final frames (statistical: photon RNG differs by
design, see sppm.cu)."""
import json
import os
import re
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gpu_pathtracer_b200 import scene as S  # noqa: E402
from gpu_pathtracer_b200.api import PathTracer  # noqa: E402

BIN = os.path.join(ROOT, "oracle", "_ref", "gpu_path_tracing_100")


def read_p3(path):
    tok = open(path).read().split()
    w, h = int(tok[1]), int(tok[2])
    return np.asarray(tok[4:4 + 3 * w * h], dtype=np.int32).reshape(h, w, 3)


def main():
    a = sys.argv[1:]
    photons = int(a[0]) if len(a) > 0 else 1_000_000
    passes = int(a[1]) if len(a) > 1 else 2
    cells = int(a[2]) if len(a) > 2 else 91
    W = int(a[3]) if len(a) > 3 else 1920
    H = int(a[4]) if len(a) > 4 else 1080
    sc = S.make_scene("C4", cells=cells, width=W, height=H, photons=photons, passes=passes)
    folder = "/tmp/gpt_scene_C4"
    master = S.write_scene_files(sc, folder, "C4")
    t0 = time.time()
    r = subprocess.run([BIN, master, os.path.join(folder, "ref_out")], capture_output=True, text=True)
    wall = time.time() - t0
    if r.returncode != 0:
        print(r.stdout[-3000:], r.stderr[-3000:])
        raise SystemExit(f"reference binary failed with {r.returncode}")
    # stage lines in order; the SPPM loop starts at the first "PPM Pass"
    body = r.stdout[r.stdout.index("PPM Pass 0"):]
    stage = {}
    for m in re.finditer(r"Done with process : (.*)\nThe process took  :\s*([0-9.]+) seconds", body):
        stage.setdefault(m.group(1).strip(), []).append(float(m.group(2)))
    tot = lambda k: float(sum(stage.get(k, [])))
    tree_keys = [k for k in stage if k not in ("Photon pass", "Ray tracing pass",
                                               "Update hit point parameters") and
                 not k.startswith("Saving") and not k.startswith("Compute ") and
                 not k.startswith("Clearing") and not k.startswith("Creating photon image")]
    ref = dict(photon_pass_s=tot("Photon pass"), eye_pass_s=tot("Ray tracing pass"),
               gather_update_s=tot("Update hit point parameters"),
               photon_compaction_and_trees_s=float(sum(tot(k) for k in tree_keys)))
    ref["per_pass_s"] = sum(ref.values()) / passes
    counts = re.findall(r"Number of surface photons = (\d+)\nNumber of volume photons  = (\d+)", body)
    ref_img = read_p3(os.path.join(folder, "ref_out.ppm"))

    pt = PathTracer(sc)
    pt.build()
    pt.sppm_init()
    ours = dict(photon_ms=0.0, eye_ms=0.0, gather_ms=0.0, rays=0)
    for p in range(passes):
        for name, mask in (("photon", 1), ("eye", 2), ("gather", 4)):
            pt.sppm_pass_stages(p, mask)
            s = pt.stats()
            ours[name + "_ms"] += s["render_ms"]
            ours["rays"] += s["rays_traced"]
        pt.sppm_update(p)
    st = pt.sppm_stats()
    img = pt.sppm_resolve(0, passes)
    ours["per_pass_ms"] = (ours["photon_ms"] + ours["eye_ms"] + ours["gather_ms"]) / passes
    u8 = pt.tonemap_u8(img).astype(np.int32)
    out = dict(scene="C4", triangles=sc.num_triangles, frame=f"{W}x{H}", photons=photons,
               passes=passes, ref=ref, ref_total_wall_s=wall, ref_photon_counts=counts[-1:] ,
               ours=ours, ours_counts=[st["n_surface"], st["n_volume"]],
               speedup_per_pass=ref["per_pass_s"] * 1e3 / ours["per_pass_ms"],
               ours_mrays=ours["rays"] / (ours["per_pass_ms"] * passes) / 1e3,
               ref_mrays_at_our_ray_count=ours["rays"] / (ref["per_pass_s"] * passes) / 1e6,
               u8_mean_ref=float(ref_img.mean()), u8_mean_ours=float(u8.mean()),
               u8_within8_frac=float((np.abs(u8 - ref_img) <= 8).mean()))
    print(json.dumps(out))
    if len(a) > 5:
        json.dump(out, open(a[5], "w"), indent=1)


if __name__ == "__main__":
    main()
