"""``tp_columnwise`` (AllGather -> GEMM) implementations (lazy exports)."""
_P = "ddlb_b200.primitives.TPColumnwise"
_LAZY = {
    "TPColumnwise": (f"{_P}.tp_columnwise", "TPColumnwise"),
    "PyTorchTPColumnwise": (f"{_P}.pytorch", "PyTorchTPColumnwise"),
    "ComputeOnlyTPColumnwise": (f"{_P}.compute_only", "ComputeOnlyTPColumnwise"),
    "FusedTPColumnwise": (f"{_P}.fused", "FusedTPColumnwise"),
    "FuserTPColumnwise": (f"{_P}.compat", "FuserTPColumnwise"),
    "TransformerEngineTPColumnwise": (f"{_P}.compat", "TransformerEngineTPColumnwise"),
    "JAXTPColumnwise": (f"{_P}.compat", "JAXTPColumnwise"),
}
__all__ = list(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        import importlib
        module, attr = _LAZY[name]
        return getattr(importlib.import_module(module), attr)
    raise AttributeError(f"module '{_P}' has no attribute {name!r}")
