"""Runs the reference's own CUDA build (oracle/_ref/gpu_path_tracing_100, sm_100, built by
`make -C oracle refcuda`) on a scene written to files, next to our path on the same scene.

Reports: the reference's "Rendering" stage seconds (its own print_end_process line), our frame
ms, Mrays/s of both (ray count = our closest-hit count, which equals the reference's
traverse_bvh(node_list[0]) count sample-for-sample), and 8-bit image agreement.

    python tools/run_reference_cuda.py C1 [spp] [json-out]
"""
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

BIN = os.path.join(ROOT, "oracle", "_ref", os.environ.get("GPT_REF_BIN", "gpu_path_tracing_100"))


def read_p3(path):
    tok = open(path).read().split()
    assert tok[0] == "P3"
    w, h = int(tok[1]), int(tok[2])
    return np.asarray(tok[4:4 + 3 * w * h], dtype=np.int32).reshape(h, w, 3)


def to_u8(img):  # save_image (src/util/image_util.h:50-72)
    return (255 * np.clip(np.sqrt(np.clip(img, 0, None)), 0, 1)).astype(np.int32)


def main():
    kind = sys.argv[1] if len(sys.argv) > 1 else "C1"
    sc = S.make_scene(kind) if kind != "small" else S.make_scene("small", n=32, textured=True)
    if len(sys.argv) > 2 and int(sys.argv[2]) > 0:
        sc.params["pathtracing_sample_size"] = int(sys.argv[2])
    folder = f"/tmp/gpt_scene_{kind}"
    t0 = time.time()
    master = S.write_scene_files(sc, folder, kind)
    t_write = time.time() - t0
    t0 = time.time()
    r = subprocess.run([BIN, master, os.path.join(folder, "ref_out")], capture_output=True,
                       text=True)
    wall = time.time() - t0
    if r.returncode != 0:
        print(r.stdout[-3000:], r.stderr[-3000:])
        raise SystemExit(f"reference binary failed with {r.returncode}")
    stages = dict((m.group(1).strip(), float(m.group(2))) for m in re.finditer(
        r"Done with process : (.*)\nThe process took  :\s*([0-9.]+) seconds", r.stdout))
    build_keys = ["Computing the world bounding box",
                  "Computing the morton code of every geometry bounding box",
                  "Sorting the objects based on morton code", "Building leaves", "Building nodes",
                  "Extracting morton codes", "Setting node relationship",
                  "Compute node bounding boxes"]
    ref_img = read_p3(os.path.join(folder, "ref_out.ppm"))

    pt = PathTracer(sc)
    st = pt.build()
    pt.render_pathtrace()
    img = pt.render_pathtrace()
    s2 = pt.stats()
    ours = to_u8(img)
    diff = np.abs(ours - ref_img)
    out = {
        "scene": kind, "triangles": sc.num_triangles, "reference_binary": os.path.basename(BIN),
        "frame": f"{sc.camera['width']}x{sc.camera['height']} @ {sc.params['pathtracing_sample_size']} spp",
        "ref_render_s": stages.get("Rendering"), "ref_total_wall_s": wall,
        "ref_main_bvh_build_s": sum(stages.get(k, 0.0) for k in build_keys),
        "ref_stage_s": {k: stages.get(k) for k in build_keys},
        "ours_render_ms": s2["render_ms"], "ours_main_bvh_build_ms": st["build_ms"],
        "rays": s2["rays_traced"],
        "ref_mrays": s2["rays_traced"] / max(stages.get("Rendering", 0.0), 1e-9) / 1e6,
        "ours_mrays": s2["rays_traced"] / s2["render_ms"] / 1e3,
        "speedup_render": stages.get("Rendering", 0.0) * 1e3 / s2["render_ms"],
        "u8_equal_frac": float((diff == 0).mean()),
        "u8_within1_frac": float((diff <= 1).mean()),
        "u8_mean_ref": float(ref_img.mean()), "u8_mean_ours": float(ours.mean()),
        "scene_write_s": t_write,
    }
    print(json.dumps(out))
    if len(sys.argv) > 3:
        json.dump(out, open(sys.argv[3], "w"), indent=1)


if __name__ == "__main__":
    main()
