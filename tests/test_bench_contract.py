"""bench.py's roofline bookkeeping, checked without a GPU: the per-stage algorithmic bytes, the dominant-kernel block
the contract asks for (bound / achieved / peak / frac / traffic) and the issue-slot view of the blending kernels."""
import importlib.util
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_roofline_block_reports_the_dominant_kernel_and_its_committed_ncu_figures():
    b = _bench()
    t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
    P, R = t["P"], t["num_rendered"]
    stage = {"preprocess": (0.17, 20), "depth_sort": (0.18, 20), "scan": (0.04, 20), "duplicate": (0.07, 20), "tile_sort": (0.25, 20),
             "tile_ranges": (0.02, 20), "render": (0.71, 20), "bwd_zero": (0.03, 20), "render_bwd": (0.80, 20), "preprocess_bwd": (0.35, 20)}
    out = b.roofline_block(stage, P, R, 1920 * 1080, int(0.77 * P), 16, 6551.0, "measured", 20)
    rl = out["roofline"]
    assert rl["kernel"] == "render_bwd_kernel" and rl["bound"] in ("hbm", "tensor") and rl["unit"] == "GB/s"
    assert abs(rl["frac"] - rl["achieved"] / rl["peak"]) < 1e-3
    assert rl["alg_bytes_per_launch"] == 40 * R + 20 * 1920 * 1080 + 44 * R
    assert rl["traffic"] == t["kernels"]["render_bwd_kernel"]                # per launch, from the committed ncu --set full capture
    iss = rl["issue_slots"]
    assert iss["warp_instructions_per_launch"] == t["warp_instructions"]["render_bwd_kernel"]
    assert 0.5 < iss["frac"] < 1.0
    assert set(out["stages"]) == set(stage)
    for name, s in out["stages"].items():
        assert s["alg_bytes"] > 0 and abs(s["gbs"] - s["alg_bytes"] / s["ms"] / 1e6) < 1.0, name
    assert "pair_rate" in out
    assert out["stages"]["tile_sort"]["alg_bytes"] == 2 * 20 * R             # two radix passes when a ranges stage was timed
    # the counting tile sort (grids of at most 8192 tiles) has no ranges stage: the library reports it as not sampled
    stage2 = dict(stage, tile_sort=(0.155, 20), tile_ranges=(-1.0, 0))
    out2 = b.roofline_block(stage2, P, R, 1920 * 1080, int(0.77 * P), 16, 6551.0, "measured", 20)
    assert "tile_ranges" not in out2["stages"] and out2["stages"]["tile_sort"]["alg_bytes"] == 16 * R


def test_roofline_block_leaves_traffic_null_for_a_workload_without_a_capture():
    b = _bench()
    stage = {"preprocess": (0.01, 3), "render": (0.05, 3), "render_bwd": (0.04, 3), "preprocess_bwd": (0.02, 3)}
    out = b.roofline_block(stage, 10000, 41000, 256 * 256, 9000, 1, 6551.0, "measured", 3)
    assert out["roofline"]["kernel"] == "render_fwd_kernel" and out["roofline"]["traffic"] is None
    assert "issue_slots" not in out["roofline"]
