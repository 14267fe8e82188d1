from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer
sc = S.make_scene("C5", n=700, width=1920, height=1080, spp=2)
pt = PathTracer(sc); pt.build(); pt.render_pathtrace(); pt.render_pathtrace()
