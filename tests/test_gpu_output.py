"""Output stage (8 f-3) and debug splats (8 f-4)."""
import numpy as np
import pytest

from gpu_pathtracer_b200 import scene as S

pytestmark = pytest.mark.gpu


def test_tonemap_and_ppm_bit_exact(tmp_path):
    """(int)(255 * clamp(0, 1, sqrt(c))) with IEEE fp32 sqrt: bit-identical to the host formula
    of save_image, including NaN / negative / > 1 inputs; the file bytes follow."""
    from gpu_pathtracer_b200.api import PathTracer
    sc = S.make_scene("small", n=8, width=32, height=16)
    pt = PathTracer(sc, device=0)
    rng = np.random.default_rng(5)
    img = (rng.random((16, 32, 3), dtype=np.float32) * 1.5).astype(np.float32)
    img[0, 0] = (np.nan, -1.0, 4.0)
    img[0, 1] = (0.0, 1.0, np.inf)
    got = pt.tonemap_u8(img)
    with np.errstate(invalid="ignore"):
        r = np.sqrt(img)
    c = np.fmax(np.float32(0), np.fmin(np.float32(1), r))  # fminf / fmaxf drop NaN
    want = (np.float32(255) * c).astype(np.int32).astype(np.uint8)
    assert np.array_equal(got, want)
    path = str(tmp_path / "img.ppm")
    pt.save_image(img, path)
    text = "P3\n32 16\n255\n" + "".join(f"{int(v)} " for v in want.reshape(-1))
    assert open(path, "rb").read() == text.encode()


def test_sppm_debug_splats():
    from gpu_pathtracer_b200.api import PathTracer
    sc = S.make_scene("C4", cells=4, width=96, height=64, photons=20000, passes=1, sphere_subdiv=2)
    pt = PathTracer(sc, device=0)
    pt.build()
    pt.sppm_init()
    pt.sppm_pass(0)
    hp = pt.sppm_export_hit_points()
    img = pt.sppm_debug_image(0)
    # Camera::compute_i / compute_j (camera.h:54-66) restated in numpy: every finite hit point
    # must have painted the pixel it projects to
    cam = sc.camera
    eye, center, up = (np.asarray(cam[k], np.float64) for k in ("eye", "center", "up"))
    W, H = cam["width"], cam["height"]
    w = (eye - center) / np.linalg.norm(eye - center)
    u = np.cross(up, w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    fovy = cam["fovy"]
    fovx = 2 * np.degrees(np.arctan(W / H * np.tan(np.radians(fovy) / 2)))
    loc = hp["loc_radius"][:, :3].astype(np.float64)
    fin = np.isfinite(loc[:, 0])
    d = loc[fin] - eye
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    jj = np.floor(W / 2 + W / 2 / np.tan(np.radians(fovx) / 2) * ((d + w) @ u)).astype(int)
    ii = np.floor(H / 2 - H / 2 / np.tan(np.radians(fovy) / 2) * ((d + w) @ v)).astype(int)
    ok = (ii >= 0) & (jj >= 0) & (ii * W + jj < W * H) & (jj < W)
    covered = img.sum(-1) > 0
    assert ok.sum() > 1000
    assert covered[ii[ok], jj[ok]].mean() > 0.98
    filt = {tuple(np.round(f, 5)) for f in hp["filter"][:, :3]}
    ys, xs = np.nonzero(covered)
    for y, x in list(zip(ys, xs))[:50]:
        assert tuple(np.round(img[y, x], 5)) in filt
    for which in (1, 2):
        ph = pt.sppm_debug_image(which)
        assert ph.shape == (64, 96, 3) and (ph.sum(-1) > 0).sum() > 50


def test_sss_debug_splat():
    from gpu_pathtracer_b200.api import PathTracer
    sc = S.make_scene("small", n=16, width=128, height=128, spp=2, level=3, sss_pts=400)
    pt = PathTracer(sc, device=0)
    pt.build()
    pts = pt.sss_export_points(first_pass=True)
    img = pt.sss_debug_image()
    assert (img.sum(-1) > 0).sum() > 50
    assert img.max() <= pts["color"].max() + 1e-6


def test_sppm_debug_splats_equal_create_point_image():
    """f-4 value parity: on the reference's own hit points and photons the three SPPM splats
    are bit-identical to the reference's create_point_image (point_operations.h:19-75) run
    with its threads in index order -- the last writer of a pixel is the highest index of the
    REFERENCE's array order (photons: the order gather_recorded_photons leaves, before the tree
    build sorts them, src/main.cu:1534-1551)."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay, available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    sc = S.make_scene("C4", cells=6, width=128, height=104, photons=60000, passes=1, sphere_subdiv=2)
    ref = RefReplay(sc)
    ref.sppm_init()
    ref.sppm_pass(0, photon_rng=0, stages=1)
    hp, ph = ref.sppm_export_hit_points(), ref.sppm_export_photons()
    pt = PathTracer(sc, device=0)
    pt.build()
    pt.sppm_init()
    pt.sppm_import_hit_points(hp["loc_radius"], hp["normal_vradius"], hp["filter"])
    pt.sppm_import_photons(ph["kind"], ph["rec"])
    for which in (0, 1, 2):
        got, want = pt.sppm_debug_image(which), ref.point_image(which)
        painted = (want.sum(-1) > 0).sum()
        assert painted > 200, (which, painted)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), which
    # several photons per pixel: the winner rule is exercised, not only the projection
    n_surface = int((ph["kind"] == 0).sum())
    assert n_surface > 2 * (ref.point_image(1).sum(-1) > 0).sum()


def test_sss_debug_splat_equals_create_point_image():
    """The SSS point splat (src/main.cu:1244-1262: after the per-object trees sorted the points
    and do_sss_first_pass coloured them) against the reference's kernel on the replay's points.
    Positions and tree order are bit-exact (test_gpu_sss.py), so exactly the same pixels are
    painted by the same points; the colours are the first pass's 2-sample Monte-Carlo estimates,
    which agree to 1e-3 except where a libm ulp flipped a decision of a path (the tolerance of
    test_gpu_sss.py::test_first_pass_colours)."""
    from gpu_pathtracer_b200.api import PathTracer
    from tests.refreplay import RefReplay, available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    sc = S.make_scene("small", n=16, width=128, height=128, spp=2, level=3, sss_pts=400)
    ref = RefReplay(sc)
    ref.sss_export_points(first_pass=True)
    want = ref.point_image(3)
    pt = PathTracer(sc, device=0)
    pt.build()
    pt.sss_export_points(first_pass=True)
    got = pt.sss_debug_image()
    assert np.array_equal(got.sum(-1) > 0, want.sum(-1) > 0)
    painted = want.sum(-1) > 0
    assert painted.sum() > 50
    close = np.isclose(got, want, rtol=1e-3, atol=1e-4).all(-1)[painted].mean()
    print(f"SSS splat: {painted.sum()} painted pixels, {close:.4f} of them within 1e-3")
    assert close >= 0.9, close
