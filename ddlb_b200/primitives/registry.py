"""Implementation registry: ``(primitive, name) -> (module, class)``.

Names match the reference's table (``ddlb/benchmark.py:43-56``). ``fused`` is
the product implementation (one sm_100a kernel per primitive). ``fuser``,
``transformer_engine`` and ``jax`` are accepted so existing configs keep
running: they are served by the same fused kernels with the overlap schedule
that matches what the named back-end would have done (see
``ddlb_b200/primitives/TPColumnwise/compat.py``).
"""
from __future__ import annotations

import importlib
from typing import Dict, Tuple

_PKG = "ddlb_b200.primitives"

REGISTRY: Dict[str, Dict[str, Tuple[str, str]]] = {
    "tp_columnwise": {
        "pytorch": (f"{_PKG}.TPColumnwise.pytorch", "PyTorchTPColumnwise"),
        "compute_only": (f"{_PKG}.TPColumnwise.compute_only", "ComputeOnlyTPColumnwise"),
        "fused": (f"{_PKG}.TPColumnwise.fused", "FusedTPColumnwise"),
        "fuser": (f"{_PKG}.TPColumnwise.compat", "FuserTPColumnwise"),
        "transformer_engine": (f"{_PKG}.TPColumnwise.compat", "TransformerEngineTPColumnwise"),
        "jax": (f"{_PKG}.TPColumnwise.compat", "JAXTPColumnwise"),
    },
    "tp_rowwise": {
        "pytorch": (f"{_PKG}.TPRowwise.pytorch", "PyTorchTPRowwise"),
        "compute_only": (f"{_PKG}.TPRowwise.compute_only", "ComputeOnlyTPRowwise"),
        "fused": (f"{_PKG}.TPRowwise.fused", "FusedTPRowwise"),
        "fuser": (f"{_PKG}.TPRowwise.compat", "FuserTPRowwise"),
        "transformer_engine": (f"{_PKG}.TPRowwise.compat", "TransformerEngineTPRowwise"),
    },
}

ALLOWED_PRIMITIVES = frozenset(REGISTRY)


def base_impl_name(impl_id: str) -> str:
    """``"compute_only_3" -> "compute_only"``; ids without ``_<index>`` pass through."""
    head, sep, tail = impl_id.rpartition("_")
    return head if sep and tail.isdigit() else impl_id


def load_impl_class(primitive: str, base_impl: str):
    if primitive not in REGISTRY:
        raise ValueError(f"Unknown primitive: {primitive}")
    table = REGISTRY[primitive]
    if base_impl not in table:
        raise ValueError(
            f"Unknown implementation '{base_impl}' for primitive '{primitive}' "
            f"(available: {sorted(table)})")
    module_path, class_name = table[base_impl]
    return getattr(importlib.import_module(module_path), class_name)
