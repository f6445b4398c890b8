"""Two spawned ranks on gloo/CPU through the public entry points: BASELINE.json config #1
(tp_columnwise m=1024 n=128 k=1024 float32, compute_only + pytorch gloo) plus tp_rowwise,
asserting validity and the CSV schema. Also: a crashing implementation must produce an
error row instead of hanging the run (the reference hangs, ddlb/benchmark.py:369)."""
import csv
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

REFERENCE_PREFIX = [
    "implementation", "mean_time (ms)", "std_time", "min_time", "max_time", "m", "n", "k",
    "dtype", "Throughput (TFLOPS)", "Throughput std (TFLOPS)", "world_size", "hostname",
    "time_measurement_backend", "barrier_at_each_iteration", "option", "valid"]


def _run(np_, cfg, tmp_path, timeout=300):
    path = tmp_path / "cfg.json"
    path.write_text(json.dumps(cfg))
    cmd = [sys.executable, "-m", "ddlb_b200.launcher", "-np", str(np_), "--cpu",
           "--timeout", str(timeout), os.path.join(ROOT, "scripts", "run_benchmark.py"), str(path)]
    proc = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout + 60)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    with open(cfg["benchmark"]["output_csv"]) as fh:
        reader = csv.DictReader(fh)
        return reader.fieldnames, list(reader)


def test_plumbing_config_two_ranks(tmp_path):
    cfg = {"benchmark": {
        "primitive": "tp_columnwise", "m": 1024, "n": 128, "k": 1024, "dtype": "float32",
        "validate": True, "num_iterations": 2, "num_warmups": 1,
        "output_csv": str(tmp_path / "out.csv"),
        "implementations": {
            "compute_only": [{"size": ["sharded", "unsharded"]}],
            "pytorch": [{"backend": "gloo", "order": ["AG_before", "AG_after"]}]}}}
    fields, rows = _run(2, cfg, tmp_path)
    assert fields[:len(REFERENCE_PREFIX)] == REFERENCE_PREFIX
    assert fields[len(REFERENCE_PREFIX):] == ["roofline_ms", "pct_roofline", "flops_per_gpu",
                                              "comm_bytes_per_gpu", "error"]
    assert len(rows) == 4
    for r in rows:
        assert r["valid"] == "True" and r["error"] == "" and r["world_size"] == "2"
        assert float(r["mean_time (ms)"]) > 0 and float(r["Throughput (TFLOPS)"]) > 0
    assert rows[0]["option"] == "size=sharded" and rows[1]["option"] == "size=unsharded"
    assert rows[2]["implementation"].startswith("pytorch (backend=gloo, order=AG_before")


def test_rowwise_two_ranks_and_error_row(tmp_path):
    cfg = {"benchmark": {
        "primitive": "tp_rowwise", "m": 256, "n": 64, "k": 512, "dtype": "float32",
        "validate": True, "num_iterations": 2, "num_warmups": 1,
        "output_csv": str(tmp_path / "out.csv"),
        "implementations": {
            "pytorch": [{"backend": "gloo"}],
            "fused": [{}],                       # needs CUDA: must fail loudly, not hang
            "compute_only": [{"size": "unsharded"}]}}}
    _, rows = _run(2, cfg, tmp_path)
    assert len(rows) == 3
    assert rows[0]["valid"] == "True"
    assert rows[1]["valid"] == "False" and "CUDA" in rows[1]["error"]
    assert rows[2]["valid"] == "True"            # the run continued after the failure


def test_resume_skips_done_rows(tmp_path):
    out = tmp_path / "out.csv"
    cfg = {"benchmark": {
        "primitive": "tp_columnwise", "m": 128, "n": 64, "k": 64, "dtype": "float32",
        "validate": True, "num_iterations": 1, "num_warmups": 0, "output_csv": str(out),
        "resume": True,
        "implementations": {"compute_only": [{"size": "sharded"}]}}}
    _run(1, cfg, tmp_path)
    cfg["benchmark"]["implementations"]["compute_only"] = [{"size": ["sharded", "unsharded"]}]
    _, rows = _run(1, cfg, tmp_path)
    assert [r["option"] for r in rows] == ["size=sharded", "size=unsharded"]
