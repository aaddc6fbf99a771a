"""CPU tests of bench.py's host-side helpers (no GPU, no timing)."""
import importlib.util
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _bench():
    spec = importlib.util.spec_from_file_location("bench_mod", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_roofline_traffic_is_found_in_committed_ncu_summaries():
    """roofline.traffic comes from the committed `ncu --set full` summaries; templated kernels are listed as
    "void name<...>" there and must still be matched."""
    b = _bench()
    for stage in ("blend_bwd", "blend_fwd", "pose_fwd", "pose_bwd", "preprocess_fwd", "preprocess_bwd"):
        t = b.ncu_traffic(stage)
        assert t is not None and t > 0, stage
    assert b.ncu_traffic("no_such_stage") is None


def test_workload_matches_baseline_config():
    b = _bench()
    assert (b.P_DEFAULT, b.W_DEFAULT, b.H_DEFAULT) == (50000, 512, 512)
    assert "512x512" in b.METRIC and "50k" in b.METRIC
