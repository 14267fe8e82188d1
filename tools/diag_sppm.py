"""Diagnostic: SPPM stage-by-stage comparison against the host replay (GPU box)."""
import sys
import numpy as np
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer
from tests.refreplay import RefReplay

sc = S.make_scene("C4", cells=6, width=96, height=64, photons=30000, passes=3, sphere_subdiv=2)
pt = PathTracer(sc, device=0); pt.build(); ref = RefReplay(sc)
pt.sppm_init(); ref.sppm_init()
for p in range(3):
    pt.sppm_pass(p)
    ref.sppm_pass(p, photon_rng=0, stages=3)
    a, b = pt.sppm_export_hit_points(), ref.sppm_export_hit_points()
    g = ref.sppm_gather_only()
    print("pass", p, "lm ours/ref", a["lm"][:, :3].mean(), b["lm"][:, :3].mean(),
          "flux", a["flux"][:, :3].mean(), g[:, :3].mean(), "count", a["flux"][:, 3].sum(), g[:, 3].sum())
    print("   direct", a["direct"][:, :3].mean(), b["direct"][:, :3].mean(),
          "filter", a["filter"][:, :3].mean(), b["filter"][:, :3].mean())
    fa, fb = np.isfinite(a["loc_radius"][:, 0]), np.isfinite(b["loc_radius"][:, 0])
    print("   finite agree", (fa == fb).mean(), "radius", a["loc_radius"][0, 3], b["loc_radius"][0, 3])
    m = fa & fb
    dl = np.abs(a["loc_radius"][m, :3] - b["loc_radius"][m, :3]).max(1)
    print("   loc maxdiff quantiles", np.quantile(dl, [0.5, 0.9, 0.99, 1.0]))
    pa, pb = pt.sppm_export_photons(), ref.sppm_export_photons()
    print("   kinds ours", np.bincount(pa["kind"], minlength=3), "ref", np.bincount(pb["kind"], minlength=3),
          "same", (pa["kind"] == pb["kind"]).mean())
    sk = (pa["kind"] == pb["kind"]) & (pa["kind"] < 2)
    dp = np.abs(pa["rec"][sk] - pb["rec"][sk])
    print("   photon rec maxdiff quantiles per field", [np.quantile(dp[:, 3*k:3*k+3].max(1), [0.5, 0.99, 1.0]).round(5).tolist() for k in range(4)])
    d = np.abs(a["lm"][:, :3] - b["lm"][:, :3]).sum(1)
    la, lb = a["lm"][:, :3].sum(1), b["lm"][:, :3].sum(1)
    print("   lm: ours0&ref>0", ((la == 0) & (lb > 0)).sum(), "ours>0&ref0", ((la > 0) & (lb == 0)).sum(),
          "both>0", ((la > 0) & (lb > 0)).sum(), "close(1e-2) among both>0",
          np.isclose(la, lb, rtol=1e-2)[(la > 0) & (lb > 0)].mean())
    z = np.where((la == 0) & (lb > 0))[0]
    print("   zero pixels rows/cols", [(int(i) // 96, int(i) % 96) for i in z[:20]])
    bad = np.argsort(-d)[:5]
    print("   worst lm pixels", bad, d[bad], a["lm"][bad, :3].sum(1), b["lm"][bad, :3].sum(1))
    pt.sppm_update(p); ref.sppm_pass(p, photon_rng=0, stages=4)
    a, b = pt.sppm_export_hit_points(), ref.sppm_export_hit_points()
    print("   indirect", a["indirect"][:, :3].mean(), b["indirect"][:, :3].mean())
