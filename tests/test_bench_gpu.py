"""bench.py itself on a small workload (a GPU is needed): the line it prints carries every key of the measurement
contract, its numbers are consistent with each other, every leg reports identical result words and bit-exact parity
samples.  (Replaces the round-1 test that only re-read a stored JSON file.)"""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SMALL = ["--genomes", "8", "--genome-len", "300000", "--reads", "60000", "--steps", "2", "--warmup", "3", "--slots", "4",
         "--sample-genomes", "2", "--sample-reads", "20000", "--oracle-htsize", "1000003", "--c4-reads", "300", "--c4-read-len", "4000",
         "--cli-genomes", "3", "--cli-reads", "30000", "--no-gather"]


def _run(cmd, timeout=900):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout[-2000:]
    return json.loads(lines[0])


def test_bench_line_contract_and_consistency():
    d = _run([sys.executable, "bench.py"] + SMALL)
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline", "table_only", "workloads", "cli_e2e"):
        assert k in d, k
    assert d["metric"] == "kmer_lookups_per_s" and d["unit"] == "lookups/s" and d["higher_is_better"] is True and d["dtype"] == "u64"
    assert d["n_gpus"] == 1 and d["steps"] == 2 and d["warmup"] == 3 and d["scaling"] == "weak" and d["vs_baseline"] is None
    lookups = d["config"]["lookups_per_step_per_gpu"]
    assert abs(d["value"] - lookups / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic", "requested_gbs", "table_probes_per_lookup"):
        assert k in r, k
    assert r["bound"] == "hbm" and abs(r["achieved"] - d["value"] * 32 / 1e9) / r["achieved"] < 1e-6 and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    # the headline is the product default (anchor index): fewer table probes than k-mers, fewer sectors than the table-only leg
    assert "anchor" in d["config"]["path"] and d["config"]["anchor_index"]["results_equal_table_only_path"] is True
    assert r["table_probes_per_lookup"] < 0.6 and d["table_only"]["roofline"]["table_probes_per_lookup"] == 1.0
    assert r["sectors_per_lookup"] < d["table_only"]["roofline"]["sectors_per_lookup"]
    e = d["e2e"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step", "host_step_ms"):
        assert k in e, k
    assert e["results_equal_resident_path"] is True and d["table_only"]["e2e"]["results_equal_resident_path"] is True
    assert e["d2h_bytes_per_step"] == 60000 * 10 and e["h2d_bytes_per_step"] > 60000 * 40
    c = d["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample", "value_1core"):
        assert k in c, k
    assert d["parity_sample"] == "bit-exact"
    for w in ("c3", "c4"):
        assert d["workloads"][w]["parity_sample"] == "bit-exact" and d["workloads"][w]["table_only"]["results_equal"] is True, w
        assert d["workloads"][w]["value"] > 0 and d["workloads"][w]["reads_classified_frac"] > 0.5
    assert 0.99 * 60000 * 248 < d["workloads"]["c3"]["lookups_per_step_per_gpu"] <= 60000 * 248      # 2 x (150 - 27 + 1) per pair
    assert "unavailable" not in d["cli_e2e"], d["cli_e2e"]
    assert d["cli_e2e"]["reads_per_s"] > 0 and d["cli_e2e"]["csv_identical_across_runs"] is True
    assert d["gpu_launches"] > 0


def test_bench_two_ranks_carry_the_sharded_leg():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = 29400 + os.getpid() % 200
    d = _run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
              "--master-port", str(port), "bench.py", "--gpus", "2"] + SMALL)
    assert d["n_gpus"] == 2 and "sharded" in d
    s = d["sharded"]
    assert s["results_equal_replicated_path"] is True and set(s["exchanges"]) == {"peer", "peer_pipelined", "nccl"}
    for x in s["exchanges"].values():
        assert x["results_equal_replicated_path"] is True and x["value"] > 0
    assert "route" in s["exchanges"]["peer"]["phases_rank0_ms"]
