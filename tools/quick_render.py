"""Scratch driver: render a config, print timing; optionally dump the image as PPM."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer

kind = sys.argv[1] if len(sys.argv) > 1 else "C2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
sc = S.make_scene(kind)
sc.params["pathtracing_sample_size"] = spp
t = time.time(); pt = PathTracer(sc); st = pt.build()
print(f"{kind}: {sc.num_triangles} tris, upload+build {time.time()-t:.2f}s, main bvh {st['build_ms']:.3f} ms, all {st['build_all_ms']:.2f} ms")
for r in range(reps):
    t = time.time(); img = pt.render_pathtrace(); wall = time.time() - t
    st = pt.stats()
    print(f"render {st['render_ms']:.1f} ms (wall {wall*1e3:.1f}), rays {st['rays_traced']/1e6:.1f} M, "
          f"{st['rays_traced']/st['render_ms']/1e3:.1f} Mrays/s, launches {st['kernel_launches']}, "
          f"mean {img.mean():.4f}, spp/s {spp*sc.params['dof_sample_size']/(st['render_ms']*1e-3):.2f}")
if len(sys.argv) > 4:
    out = (255 * np.clip(np.sqrt(np.clip(img, 0, None)), 0, 1)).astype(np.uint8)
    with open(sys.argv[4], "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (img.shape[1], img.shape[0])); f.write(out.tobytes())
