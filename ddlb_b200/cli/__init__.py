"""Command-line and dict-config front ends."""
from .benchmark import run_benchmark  # noqa: F401

__all__ = ["run_benchmark"]
