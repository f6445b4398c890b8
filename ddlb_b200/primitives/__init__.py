"""Tensor-parallel primitives (lazy exports)."""
_LAZY = {
    "TPColumnwise": ("ddlb_b200.primitives.TPColumnwise.tp_columnwise", "TPColumnwise"),
    "TPRowwise": ("ddlb_b200.primitives.TPRowwise.tp_rowwise", "TPRowwise"),
    "OptionsManager": ("ddlb_b200.primitives.options", "OptionsManager"),
    "EnvVarGuard": ("ddlb_b200.primitives.options", "EnvVarGuard"),
}
__all__ = list(_LAZY)


def __getattr__(name):
    if name in _LAZY:
        import importlib
        module, attr = _LAZY[name]
        return getattr(importlib.import_module(module), attr)
    raise AttributeError(f"module 'ddlb_b200.primitives' has no attribute {name!r}")
