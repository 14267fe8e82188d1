"""SPPM parity (SURVEY.md 8 a-15 .. a-19): the CUDA passes against the reference's own kernels
replayed on the host (oracle/ref_replay.cpp).  Tolerances are stated per test: everything that
only uses + - * / sqrt min max is compared bit for bit, anything that goes through libm
(powf, cos, exp, log ...) differs between CUDA's and glibc's implementations in the last ulp
and is compared with a relative tolerance."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import scene as S

pytestmark = pytest.mark.gpu


def _scene(**kw):
    args = dict(cells=6, width=96, height=64, photons=30000, passes=3, sphere_subdiv=2)
    args.update(kw)
    return S.make_scene("C4", **args)


@pytest.fixture(scope="module")
def pair():
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay, available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    sc = _scene()
    pt = PathTracer(sc, device=0)
    pt.build()
    ref = RefReplay(sc)
    pt.sppm_init()
    ref.sppm_init()
    return sc, pt, ref


def _close_frac(a, b, rtol, atol=1e-6):
    ok = np.isclose(a, b, rtol=rtol, atol=atol) | (np.isinf(a) & np.isinf(b) & (a == b))
    return ok.all(axis=-1).mean() if ok.ndim > 1 else ok.mean()


def test_init_hit_points(pair):
    """Eye pass (init): per-pixel RNG streams are deterministic in the reference, so hit
    points agree pixel by pixel (libm ulps only); the global mean radius is summed in the
    reference's serial order."""
    sc, pt, ref = pair
    a, b = pt.sppm_export_hit_points(), ref.sppm_export_hit_points()
    fa, fb = np.isfinite(a["loc_radius"][:, 0]), np.isfinite(b["loc_radius"][:, 0])
    assert (fa == fb).mean() >= 0.995
    m = fa & fb
    assert _close_frac(a["loc_radius"][m, :3], b["loc_radius"][m, :3], 1e-4, 1e-4) >= 0.99
    assert _close_frac(a["normal_vradius"][m, :3], b["normal_vradius"][m, :3], 1e-3, 1e-4) >= 0.99
    # every hit point carries the global mean radius (src/main.cu:1353)
    ra, rb = a["loc_radius"][:, 3], b["loc_radius"][:, 3]
    assert np.all(ra == ra[0]) and np.all(rb == rb[0])
    assert abs(ra[0] - rb[0]) <= 2e-3 * rb[0]
    assert np.allclose(a["normal_vradius"][:, 3], 2 * ra[0])
    assert _close_frac(a["direct"][m, :3], b["direct"][m, :3], 2e-3, 1e-4) >= 0.97


def test_photon_pass_one_to_one(pair):
    """Photon trace with per-photon streams (rng_mode 0): photon k of the CUDA pass is photon
    k of the reference's photon_pass launched per photon."""
    sc, pt, ref = pair
    pt.sppm_pass(0)
    ref.sppm_pass(0, photon_rng=0, stages=1)
    a, b = pt.sppm_export_photons(), ref.sppm_export_photons()
    same_kind = a["kind"] == b["kind"]
    assert same_kind.mean() >= 0.995
    st = same_kind & (a["kind"] < 2)
    assert st.sum() > 1000
    assert _close_frac(a["loc"][st], b["loc"][st], 1e-3, 1e-3) >= 0.99
    assert _close_frac(a["color"][st], b["color"][st], 1e-3, 1e-5) >= 0.99
    assert _close_frac(a["dir"][st], b["dir"][st], 1e-3, 1e-4) >= 0.99
    sa, sb = pt.sppm_stats(), ref.sppm_counts()
    assert abs(sa["n_surface"] - sb["n_surface"]) <= 0.01 * sb["n_surface"] + 5
    assert abs(sa["n_volume"] - sb["n_volume"]) <= 0.01 * sb["n_volume"] + 5


def test_gather_bit_exact_on_reference_inputs(pair):
    """Gather on the reference's own photons and hit points: photon LBVH, sphere/box descent
    order, visibility rays and the fp32 flux sums are all libm-free, so (flux, count) per hit
    point are bit-identical."""
    sc, pt, ref = pair
    ref.sppm_pass(0, photon_rng=0, stages=1)
    ph = ref.sppm_export_photons()
    hp = ref.sppm_export_hit_points()
    want = ref.sppm_gather_only()
    pt.sppm_import_hit_points(hp["loc_radius"], hp["normal_vradius"], hp["filter"])
    pt.sppm_import_photons(ph["kind"], ph["rec"])
    pt.sppm_pass_stages(0, 4)
    got = pt.sppm_export_hit_points()["flux"]
    assert np.array_equal(got[:, 3], want[:, 3]), "photon counts differ"
    assert np.array_equal(got[:, :3].view(np.uint32), want[:, :3].view(np.uint32))
    assert want[:, 3].sum() > 0


def test_eye_pass_volume_term_on_reference_photons(pair):
    """Eye pass of pass 1 against the reference's own volume photons: hit points agree pixel by
    pixel and the (prefix-weighted, order dependent) volume term agrees to libm accuracy --
    which also pins the sphere tree, its visit order and the beam kernel."""
    sc, pt, ref = pair
    ref.sppm_pass(1, photon_rng=0, stages=3)
    ph = ref.sppm_export_photons()
    hp0 = pt.sppm_export_hit_points()
    rb = ref.sppm_export_hit_points()
    # same radii as the reference so the beam radius is identical
    lr, nv = hp0["loc_radius"].copy(), hp0["normal_vradius"].copy()
    lr[:, 3], nv[:, 3] = rb["loc_radius"][:, 3], rb["normal_vradius"][:, 3]
    pt.sppm_import_hit_points(lr, nv, None)
    pt.sppm_import_photons(ph["kind"], ph["rec"])
    pt.sppm_pass_stages(1, 2)
    a = pt.sppm_export_hit_points()
    fa, fb = np.isfinite(a["loc_radius"][:, 0]), np.isfinite(rb["loc_radius"][:, 0])
    assert (fa == fb).mean() >= 0.995
    m = fa & fb
    assert _close_frac(a["loc_radius"][m, :3], rb["loc_radius"][m, :3], 1e-4, 1e-4) >= 0.99
    la, lb = a["lm"][:, :3], rb["lm"][:, :3]
    assert (lb.sum(1) > 0).sum() > 100, "scene has no volume contribution"
    assert _close_frac(la, lb, 2e-3, 1e-3 * float(lb.max())) >= 0.98
    assert abs(la.sum() - lb.sum()) <= 5e-3 * lb.sum()
    assert _close_frac(a["filter"][m, :3], rb["filter"][m, :3], 1e-3, 1e-5) >= 0.99


def test_full_frame_statistics():
    """Whole frame, 3 passes at a tiny photon budget: direct image within 0.2 %, indirect /
    global mean luminance within 6 %, radius schedule within 0.2 %, photon counts within 1 %."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay, available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    sc = _scene()
    passes = 3
    pt = PathTracer(sc, device=0)
    pt.build()
    ref = RefReplay(sc)
    pt.sppm_init()
    ref.sppm_init()
    for p in range(passes):
        pt.sppm_pass(p)
        pt.sppm_update(p)
        ref.sppm_pass(p, photon_rng=0)
        sa, sb = pt.sppm_stats(), ref.sppm_counts()
        assert abs(sa["n_surface"] - sb["n_surface"]) <= 0.01 * sb["n_surface"] + 5
        assert abs(sa["n_volume"] - sb["n_volume"]) <= 0.01 * sb["n_volume"] + 5
        assert abs(sa["mean_radius"] - sb["mean_radius"]) <= 2e-3 * sb["mean_radius"]
    # direct light is deterministic per pixel; the indirect term goes through the visibility
    # test `rec.t > |prev - x|` whose two sides are equal up to rounding (prev lies ON a
    # surface), and through the prefix-weighted volume sum: both amplify last-ulp libm
    # differences into photon-level noise, so only its statistics are compared (the same
    # stages are compared exactly on identical inputs in the tests above).
    a, b = pt.sppm_resolve(1, passes), ref.sppm_resolve(1, passes)
    assert abs(a.mean() - b.mean()) <= 2e-3 * b.mean()
    for mode in (0, 2):
        a, b = pt.sppm_resolve(mode, passes), ref.sppm_resolve(mode, passes)
        la, lb = a.mean(), b.mean()
        assert abs(la - lb) <= 0.06 * lb + 1e-6, (mode, la, lb)


def test_multi_rank_photon_split_sums_to_single_rank():
    """Surface flux and counts are linear in the photon set: two half ranges, summed, equal
    the full range (counts exactly, flux to fp32 summation order)."""
    from gpu_pathtracer_b200.api import PathTracer
    sc = _scene(fog=False)
    n = sc.params["ppm_num_photon_per_pass"]
    full = PathTracer(sc, device=0)
    full.build()
    full.sppm_init()
    full.sppm_pass(0)
    want = full.sppm_export_hit_points()["flux"]
    parts = []
    for lo, hi in ((0, n // 2), (n // 2, n)):
        pt = PathTracer(sc, device=0)
        pt.build()
        pt.sppm_init()
        pt.sppm_pass(0, lo, hi)
        parts.append(pt.sppm_export_hit_points()["flux"])
    got = parts[0] + parts[1]
    assert np.array_equal(got[:, 3], want[:, 3])
    assert np.allclose(got[:, :3], want[:, :3], rtol=1e-4, atol=1e-4)


def test_sharded_eye_pass_tiles_the_unsharded_pass():
    """r02, SURVEY 8(e) second variant: the eye pass of rank r of 2 (tile ownership only, no
    communicator: gpt_sppm_debug_eye_shard) leaves exactly the unsharded pass's hit points on its
    own 32x32 tiles and MINUS zeros elsewhere (the identity of the sum that also keeps a -0),
    so the sum of the shards -- what the all-reduce of the real thing computes -- IS the
    unsharded pass, bit for bit, volume term included."""
    from gpu_pathtracer_b200.api import PathTracer
    sc = _scene(width=128, height=96)
    W, H = sc.camera["width"], sc.camera["height"]

    def run(rank, nranks):
        pt = PathTracer(sc, device=0)
        pt.build()
        pt.sppm_init()
        pt.sppm_pass_stages(0, 1)      # photon trace + trees: identical on every context
        pt.sppm_debug_eye_shard(rank, nranks)
        pt.sppm_pass_stages(0, 2)      # the eye pass
        return pt.sppm_export_hit_points()

    def run_init(rank, nranks):
        # the init pass is sharded the same way (5 eye paths per pixel and the radius estimate);
        # without a communicator the global mean radius that follows is meaningless, so only
        # what the eye paths themselves produce is compared
        pt = PathTracer(sc, device=0)
        pt.build()
        pt.sppm_debug_eye_shard(rank, nranks)
        pt.sppm_init()
        return pt.sppm_export_hit_points()

    rows0, cols0 = np.divmod(np.arange(W * H), W)
    owner0 = ((rows0 // 32) * ((W + 31) // 32) + cols0 // 32) % 2
    init_base, init_shards = run_init(0, 1), [run_init(r, 2) for r in range(2)]
    for r, sh in enumerate(init_shards):
        mine = owner0 == r
        for k in ("loc_radius", "normal_vradius", "filter", "direct"):
            assert np.array_equal(sh[k][mine, :3].view(np.uint32), init_base[k][mine, :3].view(np.uint32)), (r, k)
        assert (sh["loc_radius"][~mine, :3].view(np.uint32) == 0x80000000).all()

    base = run(0, 1)
    shards = [run(r, 2) for r in range(2)]
    rows, cols = np.divmod(np.arange(W * H), W)
    owner = ((rows // 32) * ((W + 31) // 32) + cols // 32) % 2
    keys = (("loc_radius", 3), ("normal_vradius", 3), ("filter", 3), ("lm", 3))
    for r, sh in enumerate(shards):
        mine = owner == r
        assert mine.sum() > 1000
        for k, c in keys:
            assert np.array_equal(sh[k][mine, :c].view(np.uint32), base[k][mine, :c].view(np.uint32)), (r, k)
            assert (sh[k][~mine, :c].view(np.uint32) == 0x80000000).all(), (r, k)
        # radii are replicated state, not part of the exchange
        assert np.array_equal(sh["loc_radius"][:, 3], base["loc_radius"][:, 3])
        assert np.array_equal(sh["direct"][mine].view(np.uint32), base["direct"][mine].view(np.uint32))
    for k, c in keys:
        total = shards[0][k][:, :c] + shards[1][k][:, :c]
        assert np.array_equal(total.view(np.uint32), base[k][:, :c].view(np.uint32)), k
    assert (base["lm"][:, :3] > 0).any(), "the scene has no volume term to exchange"


def test_wavefront_photon_pass_equals_the_megakernel(monkeypatch):
    """The photon pass runs as a wavefront (k_photon_step / k_ph_trace / coherence sort); the
    one-thread-per-photon megakernel it replaced is kept as a cross-check.  Same photons, bit for
    bit, over two passes (the per-photon RNG state carries over), and the same ray count."""
    from gpu_pathtracer_b200.api import PathTracer
    sc = S.make_scene("C4", width=64, height=48)
    sc.params["ppm_num_photon_per_pass"] = 30000
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("GPT_SPPM_MEGAKERNEL", mode)
        pt = PathTracer(sc)
        pt.build()
        pt.sppm_init()
        recs = []
        for ps in range(2):
            pt.sppm_pass_stages(ps, 1, 0, 30000)
            ph = pt.sppm_export_photons()
            recs.append((ph["kind"].copy(), ph["rec"].copy(), pt.stats()["rays_traced"]))
        out[mode] = recs
    for (k0, r0, n0), (k1, r1, n1) in zip(out["0"], out["1"]):
        assert n0 == n1 and n0 > 30000
        assert np.array_equal(k0, k1)
        assert np.array_equal(r0.view(np.uint32), r1.view(np.uint32))
        assert (k0 == 1).sum() > 100 and (k0 == 0).sum() > 1000  # volume and surface photons
