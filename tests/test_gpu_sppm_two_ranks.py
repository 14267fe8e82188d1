"""Two ranks on two GPUs through the library's own NCCL communicator (gpt_comm_init): the
sharded SPPM pass (photons split by index range, volume photons all-gathered, eye pass by tiles,
hit points all-reduced, flux all-reduced) against the one-rank pass on the same scene.  Skipped
on a one-GPU box; the single-device halves of the argument are in test_gpu_sppm.py."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.multigpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, sys
import numpy as np
sys.path.insert(0, ROOT)
from gpu_pathtracer_b200 import scene as S
from gpu_pathtracer_b200.api import PathTracer
rank, nranks, idfile, out = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4]
sc = S.make_scene("C4", cells=6, width=128, height=96, photons=40000, passes=2, sphere_subdiv=2)
pt = PathTracer(sc, device=rank)
if nranks > 1:
    import time
    if rank == 0:
        uid = PathTracer.comm_unique_id()
        open(idfile + ".tmp", "wb").write(uid)
        os.rename(idfile + ".tmp", idfile)
    else:
        while not os.path.exists(idfile):
            time.sleep(0.01)
        uid = open(idfile, "rb").read()
    pt.comm_init(uid, rank, nranks)
pt.build()
pt.sppm_init()
n = sc.params["ppm_num_photon_per_pass"]
b, e = n * rank // nranks, n * (rank + 1) // nranks
for p in range(2):
    pt.sppm_pass(p, b, e)
    pt.sppm_allreduce()
    hp = pt.sppm_export_hit_points() if p == 0 else None
    if p == 0:
        first = {k: hp[k].copy() for k in ("loc_radius", "normal_vradius", "filter", "direct", "flux", "lm")}
    pt.sppm_update(p)
img = pt.sppm_resolve(0, 2)
st = pt.sppm_stats()
np.savez(out, img=img, n_volume=st["n_volume"], **first)
'''.replace("ROOT", repr(ROOT))


def _run(nranks, tmp_path):
    idf = str(tmp_path / f"id{nranks}")
    procs = []
    for r in range(nranks):
        out = str(tmp_path / f"n{nranks}_r{r}.npz")
        procs.append((subprocess.Popen([sys.executable, "-c", WORKER, str(r), str(nranks), idf, out],
                                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True), out))
    res = []
    for p, out in procs:
        so, se = p.communicate(timeout=600)
        assert p.returncode == 0, se[-3000:]
        res.append(dict(np.load(out)))
    return res


def test_two_rank_sppm_pass_equals_the_one_rank_pass(tmp_path):
    one = _run(1, tmp_path)[0]
    two = _run(2, tmp_path)
    for k in ("loc_radius", "normal_vradius", "filter", "direct"):
        # every rank holds every hit point after the exchange, bit-identical to the one-rank pass
        for r in range(2):
            assert np.array_equal(two[r][k].view(np.uint32), one[k].view(np.uint32)), (k, r)
    # the volume term: whole-pass sphere tree on every rank => the single-GPU value, exactly
    assert int(two[0]["n_volume"]) == int(one["n_volume"])
    for r in range(2):
        assert np.array_equal(two[r]["lm"].view(np.uint32), one["lm"].view(np.uint32)), r
    # surface flux: counts exactly, flux to fp32 summation order (two partial sums)
    assert np.array_equal(two[0]["flux"][:, 3], one["flux"][:, 3])
    assert np.allclose(two[0]["flux"][:, :3], one["flux"][:, :3], rtol=1e-4, atol=1e-4)
    assert np.array_equal(two[0]["img"].view(np.uint32), two[1]["img"].view(np.uint32))
    assert np.allclose(two[0]["img"], one["img"], rtol=2e-3, atol=1e-4)
