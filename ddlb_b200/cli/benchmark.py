"""Config expansion, ``run_benchmark(config)`` and the flag-based CLI.

Schema, id scheme, ``{timestamp}`` substitution, default CSV path and console
output follow the reference (``ddlb/cli/benchmark.py:85-223``); the CLI flags
are the reference's (``:229-291``) plus ``--no-validate``, ``--resume``,
``--time-measurement-backend`` and ``--no-barrier``, and ``--primitive``
accepts both primitives (the reference's parser only lists ``tp_columnwise``).
"""
from __future__ import annotations

import argparse
import itertools
from datetime import datetime
from typing import Any, Dict, List, Optional, Tuple

from ..envs import get_rank, get_world_size

REQUIRED_KEYS = ("primitive", "m", "n", "k", "dtype", "validate", "num_iterations",
                 "num_warmups", "implementations")


def _infer_scalar(value: str) -> Any:
    """``"true" -> True``, ``"12" -> 12``, ``"1.5" -> 1.5``, ``"08" -> "08"``, else str."""
    v = value.strip()
    if v.lower() in ("true", "false"):
        return v.lower() == "true"
    digits = v.lstrip("+-")
    if len(digits) > 1 and digits[0] == "0" and digits[1].isdigit():
        return v                               # "08" stays a string (not octal, not 8)
    for cast in (int, float):
        try:
            return cast(v)
        except ValueError:
            pass
    return v


def _parse_value_list(text: Optional[str]) -> Any:
    parts = [p.strip() for p in (text or "").split(",") if p.strip()]
    if not parts:
        return ""
    if len(parts) == 1:
        return _infer_scalar(parts[0])
    return [_infer_scalar(p) for p in parts]


def _parse_int_list(text: Any) -> List[int]:
    return [int(s) for s in (t.strip() for t in str(text).split(",")) if s]


def _parse_impl_spec(spec: Optional[str]) -> Tuple[str, Dict[str, Any]]:
    """``name;key=v1,v2;key2=v;flag`` -> ``(name, {key: [v1, v2], key2: v, flag: True})``."""
    if spec is None:
        raise ValueError("Empty implementation spec")
    tokens = [t for t in str(spec).split(";") if t.strip()]
    if not tokens:
        raise ValueError("Invalid implementation spec: empty")
    name = tokens[0].strip()
    options: Dict[str, Any] = {}
    for tok in tokens[1:]:
        key, sep, value = tok.partition("=")
        key = key.strip()
        if not key:
            continue
        options[key] = _parse_value_list(value.strip()) if sep else True
    return name, options


def generate_config_combinations(
        config: Dict[str, List[Dict[str, Any]]]) -> Dict[str, List[Dict[str, Any]]]:
    """Per implementation and base dict: cartesian product over list-valued options."""
    expanded: Dict[str, List[Dict[str, Any]]] = {}
    for impl_name, base_configs in config.items():
        out: List[Dict[str, Any]] = []
        for base in base_configs:
            list_keys = [k for k, v in base.items() if isinstance(v, list)]
            if not list_keys:
                out.append(base)
                continue
            for combo in itertools.product(*(base[k] for k in list_keys)):
                variant = dict(base)
                variant.update(zip(list_keys, combo))
                out.append(variant)
        expanded[impl_name] = out
    return expanded


def expand_implementations(implementations: Dict[str, List[Dict[str, Any]]]):
    """Ids ``"{name}_{i}"`` and their option dicts (reference ``:170-177``)."""
    expanded = generate_config_combinations(implementations)
    ids: List[str] = []
    options: Dict[str, Dict[str, Any]] = {}
    for impl_name, variants in expanded.items():
        for i, opts in enumerate(variants):
            impl_id = f"{impl_name}_{i}"
            ids.append(impl_id)
            options[impl_id] = {"implementation": impl_name, **opts}
    return expanded, ids, options


def resolve_output_csv(output_csv: Optional[str], primitive: str, m_list, n_list, k_list,
                       dtype: str, n_shapes: int, timestamp: Optional[str] = None) -> str:
    ts = timestamp or datetime.now().strftime("%Y%m%d_%H%M%S")
    if output_csv is not None and str(output_csv).strip():
        return str(output_csv).replace("{timestamp}", ts)
    label = f"{m_list[0]}x{k_list[0]}x{n_list[0]}" if n_shapes > 0 else "shapes"
    return f"results/{primitive}_{label}_{dtype}_{ts}.csv"


def run_benchmark(config: dict):
    """Run every implementation variant at every shape; returns the result frame."""
    import pandas as pd
    from ..benchmark import PrimitiveBenchmarkRunner

    rank, world_size = get_rank(), get_world_size()
    bench = config["benchmark"]
    missing = [key for key in REQUIRED_KEYS if key not in bench]
    if missing:
        raise KeyError(f"benchmark config is missing required keys: {missing}")
    primitive, dtype = bench["primitive"], bench["dtype"]
    as_list = lambda x: x if isinstance(x, list) else [x]
    m_list, n_list, k_list = as_list(bench["m"]), as_list(bench["n"]), as_list(bench["k"])
    shapes = list(itertools.product(m_list, n_list, k_list))
    expanded, impl_ids, impl_options = expand_implementations(bench["implementations"])

    if rank == 0:
        print(f"Running {primitive} benchmark with {world_size} processes")
        print(f"Number of shapes: {len(shapes)}")
        print("Shapes:")
        for shape in shapes:
            print(f"  {shape}")
        print("\nConfigurations:")
        for impl_name, variants in expanded.items():
            for i, opts in enumerate(variants):
                print(f"  {impl_name}_{i}: {opts}")

    csv_path = resolve_output_csv(bench.get("output_csv"), primitive, m_list, n_list, k_list,
                                  dtype, len(shapes))
    frames = []
    for (mm, nn, kk) in shapes:
        if rank == 0:
            print(f"\n--- Running shape ({mm}, {nn}, {kk}) ---")
        runner = PrimitiveBenchmarkRunner(
            primitive=primitive, m=mm, n=nn, k=kk, implementations=impl_ids, dtype=dtype,
            validate=bench["validate"], num_iterations=bench["num_iterations"],
            num_warmups=bench["num_warmups"], implementation_options=impl_options,
            output_csv=csv_path,
            time_measurement_backend=bench.get("time_measurement_backend", "cpu_clock"),
            barrier_at_each_iteration=bench.get("barrier_at_each_iteration", True),
            resume=bool(bench.get("resume", False)))
        frames.append(runner.run())

    results = pd.concat(frames, ignore_index=True) if frames else pd.DataFrame()
    if rank == 0 and len(results):
        print("\nBenchmark Results:")
        results["config"] = results["implementation"]
        cols = ["m", "n", "k", "config", "Throughput (TFLOPS)", "Throughput std (TFLOPS)",
                "mean_time (ms)", "std_time", "min_time", "max_time"]
        print(results[[c for c in cols if c in results.columns]])
        print(f"\nResults appended to {csv_path}")
    return results


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Run tensor-parallel primitive benchmarks from flags")
    p.add_argument("--primitive", required=True, choices=["tp_columnwise", "tp_rowwise"])
    p.add_argument("-m", "--m", required=True, help="comma-separated m sizes")
    p.add_argument("-n", "--n", required=True, help="comma-separated n sizes")
    p.add_argument("-k", "--k", required=True, help="comma-separated k sizes")
    p.add_argument("--dtype", default="float16")
    p.add_argument("--validate", dest="validate", action="store_true", default=True)
    p.add_argument("--no-validate", dest="validate", action="store_false")
    p.add_argument("--num-iterations", type=int, default=50)
    p.add_argument("--num-warmups", type=int, default=5)
    p.add_argument("--output-csv", type=str, default=None,
                   help="output CSV path; supports {timestamp}")
    p.add_argument("--time-measurement-backend", choices=["cpu_clock", "cuda_event"],
                   default="cpu_clock")
    p.add_argument("--no-barrier", dest="barrier", action="store_false", default=True,
                   help="do not synchronise ranks before each timed iteration")
    p.add_argument("--resume", action="store_true",
                   help="skip rows already present (without error) in the output CSV")
    p.add_argument("--impl", action="append", required=True,
                   help="name;key=value[,value];key2=value;flag  (repeatable)")
    return p


def config_from_args(args: argparse.Namespace) -> Dict[str, Any]:
    implementations: Dict[str, List[Dict[str, Any]]] = {}
    for spec in args.impl:
        name, opts = _parse_impl_spec(spec)
        implementations.setdefault(name, []).append(opts)
    return {"benchmark": {
        "primitive": args.primitive,
        "m": _parse_int_list(args.m), "n": _parse_int_list(args.n), "k": _parse_int_list(args.k),
        "dtype": args.dtype, "validate": args.validate,
        "num_iterations": int(args.num_iterations), "num_warmups": int(args.num_warmups),
        "output_csv": args.output_csv,
        "time_measurement_backend": args.time_measurement_backend,
        "barrier_at_each_iteration": args.barrier,
        "resume": args.resume,
        "implementations": implementations,
    }}


def main(argv: Optional[List[str]] = None) -> None:
    run_benchmark(config_from_args(build_parser().parse_args(argv)))


if __name__ == "__main__":
    main()
