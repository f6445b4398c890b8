"""B200-native tensor-parallel primitive library and benchmark harness.

Importing this package touches neither CUDA nor the native library: heavy
symbols resolve lazily (PEP 562), as in the reference (``ddlb/__init__.py:25-30``).
"""
__version__ = "0.2.0"

_LAZY = {
    "PrimitiveBenchmarkRunner": ("ddlb_b200.benchmark", "PrimitiveBenchmarkRunner"),
    "run_benchmark": ("ddlb_b200.cli.benchmark", "run_benchmark"),
    "Communicator": ("ddlb_b200.communicator", "Communicator"),
    "TPColumnwise": ("ddlb_b200.primitives.TPColumnwise.tp_columnwise", "TPColumnwise"),
    "TPRowwise": ("ddlb_b200.primitives.TPRowwise.tp_rowwise", "TPRowwise"),
}

__all__ = ["__version__", *_LAZY]


def __getattr__(name):
    if name in _LAZY:
        import importlib
        module, attr = _LAZY[name]
        value = getattr(importlib.import_module(module), attr)
        globals()[name] = value
        return value
    raise AttributeError(f"module 'ddlb_b200' has no attribute {name!r}")
