"""Real multi-GPU runs (cudaIpc symmetric heap, NVLink P2P, device barrier, fused kernels)
through the public entry points, one process per GPU via the in-tree launcher."""
import csv
import json
import os
import subprocess
import sys

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    import torch
    return torch.cuda.device_count()


def _launch(np_, cfg, tmp_path, timeout=600):
    cfg_path = tmp_path / "cfg.json"
    cfg_path.write_text(json.dumps(cfg))
    cmd = [sys.executable, "-m", "ddlb_b200.launcher", "-np", str(np_), "--timeout", str(timeout),
           os.path.join(ROOT, "scripts", "run_benchmark.py"), str(cfg_path)]
    proc = subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout + 60)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    with open(cfg["benchmark"]["output_csv"]) as fh:
        return list(csv.DictReader(fh))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_columnwise_fused_vs_nccl(world, tmp_path):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cfg = {"benchmark": {
        "primitive": "tp_columnwise", "m": 8192, "n": 1024, "k": 1024, "dtype": "bfloat16",
        "validate": True, "num_iterations": 10, "num_warmups": 3,
        "time_measurement_backend": "cuda_event", "output_csv": str(tmp_path / "out.csv"),
        "implementations": {
            "fused": [{"order": ["AG_before", "AG_after"]},
                      {"order": "AG_before", "algorithm": "coll_pipeline", "s": 2},
                      {"order": "AG_before", "algorithm": "p2p_pipeline"}],
            "pytorch": [{"backend": "nccl", "order": "AG_before"}]}}}
    rows = _launch(world, cfg, tmp_path)
    assert len(rows) == 5
    for r in rows:
        assert r["valid"] == "True" and r["error"] == "", r
        assert int(r["world_size"]) == world


@pytest.mark.parametrize("world", [2, 4, 8])
def test_rowwise_fused_vs_nccl(world, tmp_path):
    if _ngpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cfg = {"benchmark": {
        "primitive": "tp_rowwise", "m": 8192, "n": 1024, "k": 4096, "dtype": "float16",
        "validate": True, "num_iterations": 10, "num_warmups": 3,
        "time_measurement_backend": "cuda_event", "output_csv": str(tmp_path / "out.csv"),
        "implementations": {
            "fused": [{"algorithm": "default"}, {"algorithm": "coll_pipeline", "s": 2}],
            "pytorch": [{"backend": "nccl"}]}}}
    rows = _launch(world, cfg, tmp_path)
    assert len(rows) == 3
    for r in rows:
        assert r["valid"] == "True" and r["error"] == "", r
