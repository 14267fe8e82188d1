"""Kernel iteration helper: stand-alone trace kernel on C2 (R1 primary + R2 bounce rays)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from gpu_pathtracer_b200 import scene as S, rays as R
from gpu_pathtracer_b200.api import PathTracer

kind = sys.argv[1] if len(sys.argv) > 1 else "C2"
sc = S.make_scene(kind)
pt = PathTracer(sc)
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream); pt.set_stream(stream.cuda_stream)
pt.build()
r1 = R.camera_rays(sc)
h = pt.trace(r1)
hit = h["tri_id"] >= 0
d = r1[:, 4:7] / np.linalg.norm(r1[:, 4:7], axis=1, keepdims=True)
pts = r1[hit, 0:3] + h["t"][hit, None] * d[hit]
nrm = pt.export_triangles()["normal"][h["tri_id"][hit]]
nrm = np.where(((nrm * d[hit]).sum(1) > 0)[:, None], -nrm, nrm)
r2 = R.bounce_rays(pts.astype(np.float64), nrm.astype(np.float64), seed=7)
flush = torch.empty(128 << 20, dtype=torch.float32, device="cuda")
for name, rr in (("primary", r1), ("bounce", r2)):
    dr = torch.from_numpy(np.ascontiguousarray(rr)).cuda()
    dh = torch.empty((rr.shape[0], 8), dtype=torch.float32, device="cuda")
    ts = []
    for it in range(6):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream); pt.trace_device(dr.data_ptr(), rr.shape[0], dh.data_ptr()); b.record(stream)
        torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    c = pt.trace_count_device(dr.data_ptr(), rr.shape[0])
    ms = float(np.median(ts[2:]))
    print(f"{kind} {name}: {rr.shape[0]} rays, {ms:.3f} ms, {rr.shape[0]/ms/1e3:.1f} Mrays/s, "
          f"V_int {c['internal_visits']/rr.shape[0]:.1f} V_leaf {c['leaf_tests']/rr.shape[0]:.2f}, "
          f"checksum {int(dh[:,1].view(torch.int32).to(torch.int64).sum())} {float(dh[:,0][torch.isfinite(dh[:,0])].double().sum()):.6f}")
