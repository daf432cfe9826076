"""bench.py's self-checks that need no GPU: a main-line parity failure voids the value and sets the exit code, the parity
summary lands in the dicts the driver keeps whole (`e2e`, `roofline`), and the metric text follows the shape that ran."""
import importlib.util
import os
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def bench():
    spec = importlib.util.spec_from_file_location("coala_bench_under_test", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def _line(mismatches=0, leg_ok=True, extras=None):
    leg = dict(ok=leg_ok, seed_lists_checked=3, blocks_checked=4, mismatches=0)
    return {"value": 10.0, "parity": {"value_leg": dict(leg), "e2e_leg": dict(leg), "rows_checked": 100, "mismatches": mismatches, "ranks": 2},
            "e2e": {"value": 11.0}, "roofline": {}, "config": {}, "extras": extras or {}}


def test_clean_line_keeps_its_value_and_carries_the_summary(bench):
    out = _line(extras={"all_resident": {"value": 5.5, "binding_link_frac": 0.6123456},
                        "distribution": {"layouts": {"8x1": {"baseline": {"hit_ratio": 0.2, "value": 1.0}, "node_color": {"hit_ratio": 0.36, "value": 2.0}}}},
                        "backends": {"p2p": {"value": 3.0, "parity": {"ok": True, "mismatches": 0}}},
                        "stress": {"corrupt_rows": 0, "ok": True}})
    assert bench.finalize_parity(out) == 0
    assert out["value"] == 10.0 and out["e2e"]["value"] == 11.0
    for d in (out["e2e"], out["roofline"]):
        assert d["parity_ok"] is True and d["parity_rows_checked"] == 100 and d["parity_ranks"] == 2
        assert d["seed_lists_checked"] == 6 and d["stress_corrupt_rows"] == 0 and d["extras_parity_ok"] is True
    s = out["config"]["extras_summary"]
    assert s["all_resident.value"] == 5.5 and s["distribution.8x1.node_color.hit_ratio"] == 0.36 and s["backends.p2p.value"] == 3.0


@pytest.mark.parametrize("kw", [dict(mismatches=2), dict(leg_ok=False)])
def test_parity_failure_voids_the_value_and_the_exit_code(bench, kw):
    out = _line(**kw)
    assert bench.finalize_parity(out) != 0
    assert out["value"] is None and out["e2e"]["value"] is None
    assert out["value_voided_by_parity"] == 10.0 and out["e2e"]["parity_ok"] is False


def test_failing_extra_is_named_at_top_level_but_does_not_void_the_main_line(bench):
    out = _line(extras={"backends": {"nccl": {"value": 1.0, "parity": {"ok": False, "mismatches": 7}}}})
    assert bench.finalize_parity(out) == 0
    assert out["value"] == 10.0 and out["extras_parity_ok"] is False
    assert out["extras_parity_failed"] == ["extras.backends.nccl.parity"] and out["e2e"]["extras_parity_ok"] is False


def test_metric_text_follows_the_shape_and_fan_out(bench):
    a = types.SimpleNamespace(shape="papers100m", fan_out="10,10,10", batch_size=1024)
    m = bench.metric_string(a)
    assert "papers100m" in m and "3-layer" in m and "10,10,10" in m and "SAGE" not in m and "5,5" not in m
