"""Model-level compositions of the primitives (lazy exports)."""
_LAZY = {"TPMLP": ("ddlb_b200.models.tp_mlp", "TPMLP"),
         "LLAMA3_70B_MLP": ("ddlb_b200.models.tp_mlp", "LLAMA3_70B_MLP")}
__all__ = list(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        import importlib
        module, attr = _LAZY[name]
        return getattr(importlib.import_module(module), attr)
    raise AttributeError(f"module 'ddlb_b200.models' has no attribute {name!r}")
