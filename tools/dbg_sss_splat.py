import numpy as np
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer
from tests.refreplay import RefReplay
sc = S.make_scene("small", n=16, width=128, height=128, spp=2, level=3, sss_pts=400)
ref = RefReplay(sc)
rp = ref.sss_export_points(first_pass=True)
want = ref.point_image(3)
pt = PathTracer(sc, device=0)
pt.build()
gp = pt.sss_export_points(first_pass=True)
got = pt.sss_debug_image()
a, b = got.sum(-1) > 0, want.sum(-1) > 0
print("painted ours", a.sum(), "ref", b.sum(), "differ", (a != b).sum())
print("loc equal", np.array_equal(gp["loc"], rp["loc"]), "n", len(gp["loc"]), len(rp["loc"]))
print("color close", np.allclose(gp["color"], rp["color"], rtol=1e-3, atol=1e-4), "zero colours ours", (gp["color"].sum(-1) == 0).sum(), "ref", (rp["color"].sum(-1) == 0).sum())
ys, xs = np.nonzero(a != b)
for y, x in list(zip(ys, xs))[:8]:
    print(y, x, got[y, x], want[y, x])
both = a & b
print("max abs diff where both", np.abs(got - want)[both].max() if both.any() else None)
