"""Pure-Python tests of the config / CLI layer (no GPU, no torch.distributed)."""
import json
import re

import pytest

from ddlb_b200.cli import benchmark as cli


@pytest.mark.parametrize("text,expected", [
    ("true", True), ("False", False), ("12", 12), ("-3", -3), ("0", 0), ("1.5", 1.5),
    ("0.25", 0.25), ("08", "08"), ("nccl", "nccl"), ("ucc/tl/cuda", "ucc/tl/cuda"), (" 7 ", 7),
])
def test_infer_scalar(text, expected):
    got = cli._infer_scalar(text)
    assert got == expected and type(got) is type(expected)


def test_parse_value_list():
    assert cli._parse_value_list("a,b") == ["a", "b"]
    assert cli._parse_value_list("1,2, 3") == [1, 2, 3]
    assert cli._parse_value_list("x") == "x"
    assert cli._parse_value_list("") == ""
    assert cli._parse_value_list(None) == ""
    assert cli._parse_value_list("1,,2") == [1, 2]


def test_parse_int_list():
    assert cli._parse_int_list("1024, 8192,16384") == [1024, 8192, 16384]
    assert cli._parse_int_list(512) == [512]


def test_parse_impl_spec():
    name, opts = cli._parse_impl_spec("fuser;algorithm=coll_pipeline,p2p_pipeline;s=4;flag")
    assert name == "fuser"
    assert opts == {"algorithm": ["coll_pipeline", "p2p_pipeline"], "s": 4, "flag": True}
    assert cli._parse_impl_spec("pytorch") == ("pytorch", {})
    with pytest.raises(ValueError):
        cli._parse_impl_spec("")
    with pytest.raises(ValueError):
        cli._parse_impl_spec(None)


def test_generate_config_combinations():
    cfg = {"pytorch": [{"backend": "nccl", "order": ["AG_before", "AG_after"]}],
           "compute_only": [{"size": ["sharded", "unsharded"]}, {}],
           "fused": [{"algorithm": ["default", "coll_pipeline"], "s": [2, 4]}]}
    out = cli.generate_config_combinations(cfg)
    assert out["pytorch"] == [{"backend": "nccl", "order": "AG_before"},
                              {"backend": "nccl", "order": "AG_after"}]
    assert out["compute_only"] == [{"size": "sharded"}, {"size": "unsharded"}, {}]
    assert len(out["fused"]) == 4
    assert {"algorithm": "coll_pipeline", "s": 4} in out["fused"]
    # inputs are not mutated
    assert cfg["pytorch"][0]["order"] == ["AG_before", "AG_after"]


def test_expand_ids_and_roundtrip():
    from ddlb_b200.primitives.registry import base_impl_name
    _, ids, options = cli.expand_implementations(
        {"compute_only": [{"size": ["sharded", "unsharded"]}], "transformer_engine": [{}]})
    assert ids == ["compute_only_0", "compute_only_1", "transformer_engine_0"]
    assert options["compute_only_1"] == {"implementation": "compute_only", "size": "unsharded"}
    assert [base_impl_name(i) for i in ids] == ["compute_only", "compute_only", "transformer_engine"]
    assert base_impl_name("pytorch") == "pytorch"
    assert base_impl_name("compute_only") == "compute_only"


def test_output_csv_path():
    p = cli.resolve_output_csv("out/run_{timestamp}.csv", "tp_rowwise", [8], [16], [32],
                               "float16", 1, timestamp="20250101_000000")
    assert p == "out/run_20250101_000000.csv"
    p = cli.resolve_output_csv(None, "tp_rowwise", [8, 9], [16], [32], "float16", 2,
                               timestamp="T")
    assert p == "results/tp_rowwise_8x32x16_float16_T.csv"
    p = cli.resolve_output_csv("  ", "tp_columnwise", [8], [16], [32], "bfloat16", 0)
    assert re.fullmatch(r"results/tp_columnwise_shapes_bfloat16_\d{8}_\d{6}\.csv", p)


def test_cli_parser_to_config():
    args = cli.build_parser().parse_args([
        "--primitive", "tp_rowwise", "-m", "1024,2048", "-n", "128", "-k", "512",
        "--dtype", "bfloat16", "--no-validate", "--num-iterations", "7",
        "--impl", "pytorch;backend=nccl", "--impl", "fused;algorithm=default,coll_pipeline;s=2",
        "--impl", "pytorch;backend=gloo", "--time-measurement-backend", "cuda_event"])
    cfg = cli.config_from_args(args)["benchmark"]
    assert cfg["primitive"] == "tp_rowwise"
    assert cfg["m"] == [1024, 2048] and cfg["n"] == [128] and cfg["k"] == [512]
    assert cfg["validate"] is False and cfg["num_iterations"] == 7 and cfg["num_warmups"] == 5
    assert cfg["time_measurement_backend"] == "cuda_event"
    assert cfg["implementations"]["pytorch"] == [{"backend": "nccl"}, {"backend": "gloo"}]
    assert cfg["implementations"]["fused"] == [{"algorithm": ["default", "coll_pipeline"], "s": 2}]
    # defaults: validate on, primitive choices include both
    args = cli.build_parser().parse_args(
        ["--primitive", "tp_columnwise", "-m", "8", "-n", "8", "-k", "8", "--impl", "pytorch"])
    assert args.validate is True and args.dtype == "float16" and args.num_iterations == 50


def test_missing_required_key():
    with pytest.raises(KeyError):
        cli.run_benchmark({"benchmark": {"primitive": "tp_columnwise"}})


def test_shipped_configs_are_valid():
    import glob
    import os
    from ddlb_b200.primitives.registry import REGISTRY
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    paths = glob.glob(os.path.join(root, "scripts", "*.json"))
    assert paths
    for path in paths:
        with open(path) as fh:
            bench = json.load(fh)["benchmark"]
        for key in cli.REQUIRED_KEYS:
            assert key in bench, (path, key)
        for name in bench["implementations"]:
            assert name in REGISTRY[bench["primitive"]], (path, name)
