"""The bench line the round recorded (profiles/r01_bench_final2.json, written by `python bench.py` on a B200) carries
every key of the measurement contract, and its numbers are consistent with each other.  No GPU needed."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _line(name):
    return json.load(open(os.path.join(ROOT, "profiles", name)))


def test_recorded_bench_line_has_the_contract_keys():
    d = _line("r01_bench_final2.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["metric"] == "kmer_lookups_per_s" and d["unit"] == "lookups/s" and d["higher_is_better"] is True
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "k=31" in d["config"]["workload"] and "l2" in d["config"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in d["e2e"], k
    assert d["e2e"]["h2d_bytes_per_step"] > 4e8 and d["e2e"]["d2h_bytes_per_step"] == 1e8      # 10 M reads: 44 B in, 10 B out
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in d["roofline"], k
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in d["cpu_baseline"], k
    assert d["gpu_launches"] > 0 and not d["clocks"]["reasons"]
    assert d["parity_sample"] == "bit-exact" and d["e2e"]["results_equal_resident_path"] is True


def test_recorded_numbers_are_consistent():
    d = _line("r01_bench_final2.json")
    lookups = d["config"]["lookups_per_step_per_gpu"]
    # value = lookups per step / time per step; roofline = 32 B per lookup over the measured peak
    assert abs(d["value"] - lookups / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
    r = d["roofline"]
    assert r["bound"] == "hbm" and abs(r["achieved"] - d["value"] * 32 / 1e9) / r["achieved"] < 1e-6
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["traffic"] > 4 * 0.95 * lookups * 32          # the part moves a 128-byte line per 32-byte sector
    a = d["anchor_index"]
    assert a["results_equal_table_only_path"] is True and a["e2e"]["results_equal_table_only_path"] is True
    assert a["lookups_per_step_per_gpu"] == lookups and a["value"] > 2 * d["value"]
    assert a["sectors_per_lookup"] < 0.45 < 1.0 < d["config"]["sectors_per_lookup"]
    ref = _line("r01_bench_reference_final2.json")
    assert ref["impl"] == "reference" and ref["metric"] == d["metric"] and ref["e2e"]["h2d_bytes_per_step"] == 0
    assert ref["cpu_baseline"]["kind"] in ("reference", "port") and d["e2e"]["value"] > 100 * ref["value"]
