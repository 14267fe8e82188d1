"""How uneven is the traversal work across the C2 frame?  Primary rays of every image row
(pinhole camera_rays) traced with visit counting: node visits per ray by row.  The rows near
the horizon are where a pixel's dependent chain of rays is longest -- the critical path of a
tile-sharded frame.  Usage: PYTHONPATH=. python tools/row_steps.py"""
import json

import numpy as np
import torch

from gpu_pathtracer_b200 import rays as R
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

sc = S.make_scene("C2", spp=1)
pt = PathTracer(sc)
pt.build()
H, W = sc.camera["height"], sc.camera["width"]
r = torch.from_numpy(R.camera_rays(sc)).cuda().contiguous()
per_row = []
for y in range(0, H, 8):  # bands of 8 rows
    band = r[y * W:(y + 8) * W]
    c = pt.trace_count_device(band.data_ptr(), band.shape[0])
    per_row.append((c["internal_visits"] + c["leaf_tests"]) / band.shape[0])
per_row = np.array(per_row)
allc = pt.trace_count_device(r.data_ptr(), r.shape[0])
mean = (allc["internal_visits"] + allc["leaf_tests"]) / r.shape[0]
print(json.dumps({"primary_rays": int(r.shape[0]), "mean_steps_per_ray": round(float(mean), 1),
                  "band_of_8_rows_max": round(float(per_row.max()), 1),
                  "band_at_row": int(per_row.argmax() * 8),
                  "bands_over_2x_mean": int((per_row > 2 * mean).sum()), "bands": len(per_row),
                  "p50": round(float(np.median(per_row)), 1),
                  "p90": round(float(np.percentile(per_row, 90)), 1),
                  "p99": round(float(np.percentile(per_row, 99)), 1)}))
