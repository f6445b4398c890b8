"""``tp_rowwise`` (GEMM -> ReduceScatter) implementations (lazy exports)."""
_P = "ddlb_b200.primitives.TPRowwise"
_LAZY = {
    "TPRowwise": (f"{_P}.tp_rowwise", "TPRowwise"),
    "PyTorchTPRowwise": (f"{_P}.pytorch", "PyTorchTPRowwise"),
    "ComputeOnlyTPRowwise": (f"{_P}.compute_only", "ComputeOnlyTPRowwise"),
    "FusedTPRowwise": (f"{_P}.fused", "FusedTPRowwise"),
    "FuserTPRowwise": (f"{_P}.compat", "FuserTPRowwise"),
    "TransformerEngineTPRowwise": (f"{_P}.compat", "TransformerEngineTPRowwise"),
}
__all__ = list(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        import importlib
        module, attr = _LAZY[name]
        return getattr(importlib.import_module(module), attr)
    raise AttributeError(f"module '{_P}' has no attribute {name!r}")
