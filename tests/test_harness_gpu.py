"""The harness on one GPU: fused product path (world=1 -> pure tcgen05 GEMM), library
baselines, CUDA-event timing, CSV rows."""
import csv
import os

import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("primitive,impls", [
    ("tp_columnwise", {"fused": [{"order": ["AG_before", "AG_after"], "tile": ["default", "wide"]}],
                       "fuser": [{"algorithm": "coll_pipeline", "s": 2}],
                       "transformer_engine": [{}], "jax": [{}],
                       "pytorch": [{"backend": "nccl"}], "compute_only": [{"size": "unsharded"}]}),
    ("tp_rowwise", {"fused": [{"algorithm": ["default", "coll_pipeline"]}, {"tile": "wide"}],
                    "fuser": [{"algorithm": "p2p_pipeline"}], "transformer_engine": [{}],
                    "pytorch": [{"backend": "nccl"}], "compute_only": [{}]}),
])
def test_run_benchmark_single_gpu(tmp_path, monkeypatch, primitive, impls):
    from ddlb_b200.cli import run_benchmark
    for key in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(key, raising=False)
    monkeypatch.setenv("DDLB_MASTER_PORT", "23456")
    out = tmp_path / "r_{timestamp}.csv"
    cfg = {"benchmark": {
        "primitive": primitive, "m": 1024, "n": 512, "k": 512, "dtype": "bfloat16",
        "validate": True, "num_iterations": 3, "num_warmups": 1,
        "time_measurement_backend": "cuda_event", "output_csv": str(out),
        "implementations": impls}}
    df = run_benchmark(cfg)
    from ddlb_b200.cli.benchmark import expand_implementations
    n_rows = len(expand_implementations(impls)[1])
    assert len(df) == n_rows
    assert df["valid"].all(), df[["implementation", "valid", "error"]]
    assert (df["mean_time (ms)"] > 0).all()
    files = [f for f in os.listdir(tmp_path) if f.endswith(".csv")]
    assert len(files) == 1
    with open(tmp_path / files[0]) as fh:
        rows = list(csv.DictReader(fh))
    assert len(rows) == n_rows
    assert all(r["error"] == "" for r in rows)
