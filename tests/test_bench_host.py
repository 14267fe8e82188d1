"""Host-side logic of bench.py that needs no GPU: the parity block's arithmetic and verdict, the
capture files behind roofline.traffic / roofline.ncu, the reference window (host replay of the
reference, bounded), and the schema of the --impl reference line."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _windows(rng, rows=6, cols=(8, 40), h=32, w=48):
    ref = rng.random((h, w, 3), dtype=np.float32) + 0.1
    wins = [(r0, r0 + 2, cols[0], cols[1], ref[r0:r0 + 2, cols[0]:cols[1]].copy()) for r0 in range(4, 4 + rows, 2)]
    return ref, wins


def test_parity_block_accepts_equal_frames_and_rejects_biased_ones():
    rng = np.random.default_rng(7)
    ref, wins = _windows(rng)
    p = bench.parity_block(ref.copy(), wins)
    assert p["ok"] and p["frac_within_1e-3"] == 1.0 and p["frac_bit_equal"] == 1.0 and p["rmse"] == 0.0
    # ulp-level noise everywhere: still inside 1e-3 relative
    noisy = ref * (1.0 + 1e-5 * rng.standard_normal(ref.shape).astype(np.float32))
    assert bench.parity_block(noisy, wins)["ok"]
    # 10 % of the window re-drawn (unbiased): allowed, that is what a shifted RNG stream does
    redrawn = ref.copy()
    mask = rng.random(ref.shape[:2]) < 0.10
    redrawn[mask] = ref[mask] * (1.0 + 0.2 * rng.standard_normal((mask.sum(), 3)).astype(np.float32))
    q = bench.parity_block(redrawn, wins)
    assert q["ok"] and 0.8 < q["frac_within_1e-3"] < 0.97
    # a 2 % brightness bias must fail on the mean-luminance bound, NaN on the finite check
    assert not bench.parity_block(ref * 1.02, wins)["ok"]
    bad = ref.copy()
    bad[5, 10, 0] = np.nan
    assert not bench.parity_block(bad, wins)["ok"]
    # pixels outside the windows do not matter
    far = ref.copy()
    far[20:, :, :] = 99.0
    assert bench.parity_block(far, wins)["ok"]


def test_capture_files_behind_the_roofline_block_parse():
    cap, path = bench.newest_traffic_capture("C2")
    assert cap is not None and path.startswith("profiles/") and path.endswith("_trace_traffic_C2.json")
    assert 10 < cap["dram_bytes_per_ray"] < 5000 and cap["l2_bytes_per_ray"] > cap["dram_bytes_per_ray"]
    assert bench.newest_traffic_capture("no_such_workload") == (None, None)
    c = bench.newest_trace_capture_counters()
    assert c is not None and c["source"].endswith("_k_pt_trace_raw.csv")
    assert 0.0 < c["issue_slots_busy_frac"] <= 1.0 and 1.0 <= c["lanes_per_instruction"] <= 32.0


def test_reference_window_keeps_its_pixels():
    from tests.refreplay import available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    sc = bench.make_workload("small", 0)
    r = bench.reference_window(sc, 0.2, 2)
    assert r["rays"] > 0 and r["mrays"] > 0 and 0 < r["frame_fraction"] <= 1
    H, W = sc.camera["height"], sc.camera["width"]
    for r0, r1, c0, c1, px in r["windows"]:
        assert 0 <= r0 < r1 <= H and 0 <= c0 < c1 <= W and px.shape == (r1 - r0, c1 - c0, 3)
        assert np.isfinite(px).all() and px.max() > 0
    # the GPU frame is not available here: the replay against itself is the trivial parity
    frame = np.zeros((H, W, 3), np.float32)
    for r0, r1, c0, c1, px in r["windows"]:
        frame[r0:r1, c0:c1] = px
    assert bench.parity_block(frame, r["windows"])["ok"]


def test_reference_arm_prints_one_contract_line():
    from tests.refreplay import available
    if not available():
        pytest.skip("oracle/_ref/libref_replay.so not available")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload",
                          "small", "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better",
              "scaling", "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["cpu_baseline"]["kind"] == "reference"
    assert d["cpu_baseline"]["cores"] >= 1 and 0 < d["cpu_baseline"]["frame_fraction"] <= 1
