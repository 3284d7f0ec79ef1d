"""CPU checks of bench.py's record assembly: the per-kernel table and `roofline` object built from the library's trace,
and that both arms (ours / --impl reference) describe a workload with the same `config`."""
import os
import sys
import types

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _env(int8_peak=2900.0, prf_peak=57.0):
    env = types.SimpleNamespace()
    env.peaks, env.peak_src = {"hbm_gbs": 6552.0, "bf16_tflops": 1613.7, "bf16_tflops_sustained": 1373.2}, "measured"
    env.int8_peak, env.prf_peak = int8_peak, prf_peak
    return env


def test_kernel_table_and_rooflines():
    steps = 4
    trace = {
        "limb_gemm": {"n": 8, "ms": 30.4, "bytes": 8 * 4.3e9, "aes_blocks": 0.0, "tensor_ops": 8 * 1.0e13},
        "adder_level": {"n": 48, "ms": 17.6, "bytes": 48 * 0.77e9, "aes_blocks": 48 * 1.56e7, "tensor_ops": 0.0},
        "ew_binary3": {"n": 8, "ms": 2.9, "bytes": 8 * 2.4e9, "aes_blocks": 0.0, "tensor_ops": 0.0},
        "k_fill": {"n": 4, "ms": 0.1, "bytes": 0.0, "aes_blocks": 0.0, "tensor_ops": 0.0},
    }
    rows, per_step = bench.kernel_table(trace, steps)
    assert [r["kernel"] for r in rows] == ["limb_gemm", "adder_level", "ew_binary3", "k_fill"]      # by time
    assert per_step == pytest.approx(sum(v["ms"] for v in trace.values()) / steps)
    assert rows[0]["launches_per_step"] == 2 and rows[0]["int8_TOPs"] == pytest.approx(8e13 / 30.4e-3 / 1e12)
    assert rows[1]["G_aes_blocks_per_s"] == pytest.approx(48 * 1.56e7 / 17.6e-3 / 1e9)
    assert rows[2]["GBps"] == pytest.approx(8 * 2.4e9 / 2.9e-3 / 1e9) and rows[3]["GBps"] is None
    assert sum(r["share"] for r in rows) == pytest.approx(1.0)

    tensor = bench.roofline_from(_env(), rows, per_step, traffic=9.06e9)
    assert tensor["bound"] == "tensor" and tensor["unit"].startswith("TOP/s") and tensor["traffic"] == 9.06e9
    assert tensor["frac"] == pytest.approx(rows[0]["int8_TOPs"] / 2900.0)
    assert tensor["avg_launch_ms"] == pytest.approx(30.4 / 8)
    # without a probe the proxy peak is 2 x bf16 sustained and the source says so
    proxy = bench.roofline_from(_env(int8_peak=None), rows, per_step)
    assert proxy["peak"] == pytest.approx(2 * 1373.2) and "bf16" in proxy["peak_source"]

    prf_rows = [rows[1], rows[2]]
    prf = bench.roofline_from(_env(), prf_rows, per_step)
    assert prf["bound"] == "hbm" and prf["unit"] == "GB/s" and prf["frac"] == pytest.approx(prf_rows[0]["GBps"] / 6552.0)
    assert prf["prf"]["frac"] == pytest.approx(prf_rows[0]["G_aes_blocks_per_s"] / 57.0)
    assert len(bench.trim(rows, frac=0.5)) == 1 and len(bench.trim(rows, frac=2.0, cap=3)) == 3


def test_both_arms_name_the_same_config():
    names = bench.workload_names(65536, 1000000, 4096)
    assert names["matmul"] == "replicated_ring128_matmul_4096x4096x4096_truncpr40"
    assert "batch65536" in names["mlp"] and "batch1000000" in names["logreg"]
    # the dictionaries the two arms put under `config` are built by the same functions from the same arguments
    assert bench.matmul_config(names["matmul"]) == bench.matmul_config(names["matmul"])
    assert bench.predictor_config(names["mlp"], 4) == bench.predictor_config(names["mlp"], 4)
    assert set(bench.matmul_config(names["matmul"])) == {"workload", "fixed", "parties", "l2"}
    assert bench.ring_ops_per_step(4096, 4096, 4096) == 3 * 2 * 2 * 4096 ** 3


@pytest.mark.parametrize("name", ["r2_bench_n1_steps20.json", "r2_bench_n8.json", "r2_bench_reference_arm.json"])
def test_committed_records_carry_the_contract_fields(name):
    """The records under profiles/ are lines bench.py printed on B200 boxes: every key the bench contract names is
    there, on the top-level (matmul) record and on the MLP / LogReg records under `workloads`."""
    import json
    rec = json.load(open(os.path.join(ROOT, "profiles", name)))
    base = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
            "vs_baseline", "dtype", "data", "config", "e2e"}
    if rec["n_gpus"] == 1 or rec.get("impl") == "reference":       # the CPU baseline is timed at N = 1 only
        base.add("cpu_baseline")
    assert base <= set(rec) and rec["vs_baseline"] is None and "workload" in rec["config"] and "model" not in rec["config"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(rec["e2e"])
    if "cpu_baseline" in base:
        assert {"value", "unit", "cores", "kind", "sample"} <= set(rec["cpu_baseline"]) and rec["cpu_baseline"]["kind"] == "port"
    if rec.get("impl") == "reference":
        assert rec["e2e"]["h2d_bytes_per_step"] == 0 and rec["e2e"]["value"] == rec["value"]
        return
    records = [rec] + [rec["workloads"][k] for k in ("mlp", "logreg")]
    for r in records:
        assert r["gpu_launches"] > 0 and r["e2e"]["h2d_bytes_per_step"] > 0 and r["e2e"]["value"] != r["value"]
        assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(r["clocks"])
        roof = r["roofline"]
        assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(roof) and roof["bound"] in ("hbm", "tensor")
        assert roof["frac"] == pytest.approx(roof["achieved"] / roof["peak"], rel=1e-3)
        assert not set(r["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert rec["n_gpus"] in (1, 8) and (rec["n_gpus"] == 1 or "party_per_gpu" in rec)
