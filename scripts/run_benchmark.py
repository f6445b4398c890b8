#!/usr/bin/env python3
"""Run a benchmark described by a JSON file (default ``scripts/config.json``).

    python -m ddlb_b200.launcher -np 8 scripts/run_benchmark.py scripts/config.json

Same role as the reference's ``scripts/run_benchmark.py:9-30``; the parent never
touches CUDA (each implementation runs in a spawned child).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main() -> None:
    parser = argparse.ArgumentParser(description="Run tensor-parallel primitive benchmarks")
    parser.add_argument("config_path", nargs="?",
                        default=os.path.join(os.path.dirname(__file__), "config.json"))
    args = parser.parse_args()
    try:
        with open(args.config_path) as fh:
            config = json.load(fh)
    except Exception as exc:
        print(f"Failed to load config from {args.config_path}: {exc}")
        sys.exit(1)
    from ddlb_b200.cli import run_benchmark
    run_benchmark(config)


if __name__ == "__main__":
    main()
