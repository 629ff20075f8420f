"""CPU checks of the measurement bookkeeping: bench.py's helpers and the committed bench records under profiles/ (no GPU, no
oracle run).  The records are what BASELINE.md section 4 quotes; the contract keys are those the driver reads."""
import glob
import importlib.util
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_ncu_traffic_reads_the_dominant_launch():
    """several kernels match 'blend_backward' (the flagged-tile pass is a 4 us launch): the longest launch's DRAM bytes count"""
    b = load_bench()
    bwd, src = b.ncu_traffic("blend_backward")
    fwd, _ = b.ncu_traffic("blend_forward")
    assert src and src.startswith("profiles/") and os.path.exists(os.path.join(ROOT, src))
    assert 5e7 < bwd < 5e8 and 2e7 < fwd < 5e8


def test_committed_bench_records_carry_the_contract():
    need = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
            "dtype", "data", "config", "e2e", "clocks", "gpu_launches"}
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r2_bench_[1248]gpu.json")))
    assert len(files) == 4
    values = {}
    for f in files:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        assert need <= set(d), (f, need - set(d))
        assert d["unit"] == "ms/frame" and d["higher_is_better"] is False and d["vs_baseline"] is None and d["dtype"] == "f32"
        assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"])
        assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
        assert d["e2e"]["value"] > d["value"]                       # the host round trip costs something
        assert d["gpu_launches"] == d["steps"] * 15                  # 15 kernels per fwd+bwd frame at C2 (profiles/r2e_launches_C2.md)
        assert not set(d["clocks"].get("reasons", [])) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
        assert d["config"]["c5"].get("replicas_bit_equal", True) is True
        values[d["n_gpus"]] = d["value"]
        if d["n_gpus"] == 1:
            assert {"roofline", "cpu_baseline"} <= set(d)
            r = d["roofline"]
            assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r)
            assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
            assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["cores"] >= 1
            p = d["parity"]
            assert p["keys_equal"] and p["lists_equal"] and p["ranges_equal"] and p["last_equal"]
            assert p["image_max_abs"] <= 1e-4 and p["grad_n_fail"] == 0
    assert values[1] > values[2] > values[4] > values[8]            # aggregate ms/frame falls with the GPU count


def test_render_only_records():
    for n in (1, 2, 4, 8):
        d = json.loads(open(os.path.join(ROOT, "profiles", f"r2_bench_c4_render_{n}gpu.json")).read().strip().splitlines()[-1])
        assert d["n_gpus"] == n and d["unit"] == "frames/s" and d["higher_is_better"] is True
        assert d["e2e"]["d2h_bytes_per_step"] == 3840 * 2160 * 4
        assert d["value"] > 600.0 * n
