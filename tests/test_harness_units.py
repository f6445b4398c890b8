"""CPU unit tests of harness helpers: result-row labels, CSV append/resume bookkeeping,
roofline arithmetic, headless plotting, registry, launcher failure handling."""
import csv
import math
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_describe_implementation_and_option_merge():
    from ddlb_b200.benchmark import _merged_options, describe_implementation

    class Impl:
        DEFAULT_OPTIONS = {"backend": "nccl", "order": "AG_before", "size": "sharded"}

    opts = _merged_options(Impl, {"implementation": "x", "order": "AG_after", "bogus": 1})
    assert opts == {"backend": "nccl", "order": "AG_after", "size": "sharded"}   # unknown keys dropped
    label, option = describe_implementation(Impl, "pytorch", opts)
    assert label == "pytorch (backend=nccl, order=AG_after)"                     # size not in the label
    assert option == "backend=nccl, order=AG_after, size=sharded"                # but kept in option

    class NoOpts:
        DEFAULT_OPTIONS = {}
    assert describe_implementation(NoOpts, "jax", {}) == ("jax", "")


def test_append_row_csv_header_once_and_resume_keys(tmp_path):
    from ddlb_b200.benchmark import _empty_row, _rows_already_done, append_row_csv
    path = str(tmp_path / "sub" / "out.csv")
    ok = _empty_row("a (x=1)", "x=1", 8, 16, 32, "float16", "cuda_event", True, True, "")
    ok["valid"] = True
    bad = _empty_row("b", "", 8, 16, 32, "float16", "cuda_event", True, True, "RuntimeError: boom")
    append_row_csv(path, ok)
    append_row_csv(path, bad)
    with open(path) as fh:
        lines = fh.read().splitlines()
    assert lines[0].startswith("implementation,mean_time (ms),std_time,min_time,max_time,m,n,k,dtype,")
    assert sum(1 for l in lines if l.startswith("implementation,")) == 1
    with open(path) as fh:
        rows = list(csv.DictReader(fh))
    assert [r["implementation"] for r in rows] == ["a (x=1)", "b"]
    done = _rows_already_done(path)
    assert ("a (x=1)", "x=1", "8", "16", "32", "float16", "1") in done
    assert len(done) == 1                                   # rows with an error are retried
    assert _rows_already_done(None) == set() and _rows_already_done(str(tmp_path / "nope")) == set()


def test_runner_rejects_bad_arguments():
    from ddlb_b200.benchmark import PrimitiveBenchmarkRunner
    with pytest.raises(ValueError):
        PrimitiveBenchmarkRunner("tp_diagonal", 8, 8, 8, [])
    with pytest.raises(ValueError):
        PrimitiveBenchmarkRunner("tp_rowwise", 8, 8, 8, [], time_measurement_backend="sundial")


def test_registry():
    from ddlb_b200.primitives.registry import ALLOWED_PRIMITIVES, REGISTRY, load_impl_class
    assert ALLOWED_PRIMITIVES == {"tp_columnwise", "tp_rowwise"}
    # every name of the reference's table is served
    assert {"pytorch", "compute_only", "fuser", "transformer_engine", "jax"} <= set(REGISTRY["tp_columnwise"])
    assert {"pytorch", "fuser", "transformer_engine"} <= set(REGISTRY["tp_rowwise"])
    for prim, table in REGISTRY.items():
        for name in table:
            cls = load_impl_class(prim, name)
            assert isinstance(cls.DEFAULT_OPTIONS, dict)
    with pytest.raises(ValueError, match="Unknown implementation"):
        load_impl_class("tp_rowwise", "jax")
    with pytest.raises(ValueError, match="Unknown primitive"):
        load_impl_class("tp_x", "pytorch")


def test_compat_option_tables_match_reference_names():
    from ddlb_b200.primitives.TPColumnwise.compat import FuserTPColumnwise
    from ddlb_b200.primitives.TPRowwise.compat import FuserTPRowwise
    assert list(FuserTPColumnwise.DEFAULT_OPTIONS) == [
        "backend", "order", "algorithm", "s", "offset_stream_indexing_by_rank",
        "multicast_protocol", "inter_stream_synchronization"]
    assert "order" not in FuserTPRowwise.DEFAULT_OPTIONS
    assert FuserTPColumnwise.ALLOWED_VALUES["algorithm"] == ["default", "coll_pipeline", "p2p_pipeline"]
    assert FuserTPColumnwise._translate(
        {**FuserTPColumnwise.DEFAULT_OPTIONS, "algorithm": "coll_pipeline", "s": 4}) == {
        "order": "AG_before", "algorithm": "coll_pipeline", "s": 4, "rank_rotate": True,
        "transport": "auto"}
    # nvFuser's symmetric-memory backend: the multicast protocol selects the data path, and the
    # un-rotated stream order the un-rotated shard order
    cuda = {**FuserTPColumnwise.DEFAULT_OPTIONS, "backend": "cuda", "multicast_protocol": "multimem",
            "offset_stream_indexing_by_rank": False}
    assert FuserTPColumnwise._translate(cuda)["transport"] == "multimem"
    assert FuserTPColumnwise._translate(cuda)["rank_rotate"] is False
    assert FuserTPColumnwise._translate({**cuda, "multicast_protocol": "memcpy"})["transport"] == "push"
    assert FuserTPRowwise._translate({**FuserTPRowwise.DEFAULT_OPTIONS, "backend": "cuda",
                                      "multicast_protocol": "multimem"})["transport"] == "multimem"


def test_compat_rows_are_labelled_as_served_by_the_fused_kernel():
    from ddlb_b200.benchmark import _merged_options, describe_implementation
    from ddlb_b200.primitives.TPColumnwise.compat import (FuserTPColumnwise,
                                                           TransformerEngineTPColumnwise)
    from ddlb_b200.primitives.TPColumnwise.fused import FusedTPColumnwise
    opts = _merged_options(FuserTPColumnwise, {"backend": "nccl", "algorithm": "coll_pipeline", "s": 4,
                                               "inter_stream_synchronization": True})
    label, option = describe_implementation(FuserTPColumnwise, "fuser", opts)
    assert label.startswith("fused[as fuser] (") and "backend" not in label
    assert "algorithm=coll_pipeline" in label and "s=4" in label
    assert "ignored: " in option and "backend=nccl" in option.split("ignored: ")[1]
    assert "inter_stream_synchronization=True" in option.split("ignored: ")[1]
    label, option = describe_implementation(
        FuserTPColumnwise, "fuser", _merged_options(FuserTPColumnwise, {"backend": "cuda",
                                                                        "multicast_protocol": "multimem"}))
    assert "multicast_protocol=multimem" in label and "backend=cuda" in label
    label, _ = describe_implementation(TransformerEngineTPColumnwise, "transformer_engine", {})
    assert label == "fused[as transformer_engine]"
    label, option = describe_implementation(FusedTPColumnwise, "fused",
                                            _merged_options(FusedTPColumnwise, {}))
    assert label.startswith("fused (") and "ignored" not in option


def test_roofline_arithmetic(monkeypatch, tmp_path):
    from ddlb_b200.utils import roofline
    peaks = tmp_path / "peaks.json"
    peaks.write_text('{"bf16_tflops": 1000.0, "bf16_tflops_sustained": 800.0, "hbm_gbs": 6000.0}')
    monkeypatch.setenv("DDLB_MEASURED_PEAKS", str(peaks))
    roofline.measured_peaks.cache_clear()
    try:
        assert roofline.measured_peaks()["source"] == "measured"
        # compute bound: 2e12 FLOP at 1000 TF = 2 ms
        assert math.isclose(roofline.roofline_ms(2e12, 0.0, "bfloat16"), 2.0)
        # comm bound: 9e9 B at 900 GB/s = 10 ms
        assert math.isclose(roofline.roofline_ms(1e9, 9e9, "float16"), 10.0)
        assert math.isclose(roofline.roofline_ms(1e9, 7.7e9, "float16", link_gbs=770.0), 10.0)
        assert roofline.gemm_peak_tflops("float32") < roofline.gemm_peak_tflops("bfloat16")
        monkeypatch.setenv("DDLB_MEASURED_PEAKS", str(tmp_path / "missing.json"))
        roofline.measured_peaks.cache_clear()
        assert roofline.measured_peaks()["source"] == "fallback"
    finally:
        roofline.measured_peaks.cache_clear()


def test_bar_chart_writes_a_file(tmp_path):
    from ddlb_b200.utils.plot import bar_chart
    path = bar_chart(["pytorch (nccl)", "fused <default>"], [0.28, 0.16], [0.03, float("nan")],
                     "TP_COLUMNWISE Benchmark", "Time (ms)", str(tmp_path / "chart"))
    assert os.path.exists(path) and os.path.getsize(path) > 200
    if path.endswith(".svg"):
        text = open(path).read()
        assert "<svg" in text and "fused &lt;default&gt;" in text and text.count("<rect") >= 3


def test_launcher_propagates_failure_and_stops_other_ranks(tmp_path):
    script = tmp_path / "fail.py"
    script.write_text("import os, sys, time\n"
                      "if os.environ['RANK'] == '1': sys.exit(3)\n"
                      "time.sleep(60)\n")
    t0 = __import__("time").time()
    proc = subprocess.run([sys.executable, "-m", "ddlb_b200.launcher", "-np", "2", "--cpu", str(script)],
                          cwd=ROOT, capture_output=True, text=True, timeout=50)
    assert proc.returncode == 3
    assert __import__("time").time() - t0 < 40          # rank 0 was stopped, not waited for


def test_fused_tile_option_is_validated():
    from ddlb_b200.primitives.TPColumnwise.fused import FusedTPColumnwise
    from ddlb_b200.primitives.TPRowwise.fused import FusedTPRowwise
    for cls in (FusedTPColumnwise, FusedTPRowwise):
        assert cls.DEFAULT_OPTIONS["tile"] == "default"
        assert cls.ALLOWED_VALUES["tile"] == ["default", "wide"]


def test_launcher_sets_rank_environment(tmp_path):
    script = tmp_path / "env.py"
    script.write_text("import os\n"
                      "print(os.environ['RANK'], os.environ['WORLD_SIZE'], os.environ['LOCAL_RANK'],\n"
                      "      os.environ['DDLB_DEVICE'], bool(os.environ.get('DDLB_RDZV_PREFIX')))\n")
    proc = subprocess.run([sys.executable, "-m", "ddlb_b200.launcher", "-np", "3", "--cpu",
                           "--all-output", str(script)], cwd=ROOT, capture_output=True, text=True, timeout=60)
    assert proc.returncode == 0, proc.stderr
    assert sorted(proc.stdout.split("\n")[:3]) == ["0 3 0 cpu True", "1 3 1 cpu True", "2 3 2 cpu True"]


def test_native_library_loads_without_gpu_and_reports_errors():
    import ctypes
    from ddlb_b200 import _native as N
    lib = N.lib()
    assert lib.ddlb_version() >= 100
    import json
    topo = json.loads(N.topology_json())
    assert "device_count" in topo and isinstance(topo["devices"], list)
    import torch
    if not torch.cuda.is_available():
        cfg = N.TpGemmConfig()
        cfg.mode, cfg.dtype, cfg.rank, cfg.world, cfg.m, cfg.n, cfg.k = 0, 1, 0, 1, 128, 128, 64
        handle = ctypes.c_void_p()
        assert lib.ddlb_tp_gemm_create(ctypes.byref(cfg), ctypes.byref(handle)) != 0
        assert lib.ddlb_last_error()                    # a message, not a crash
