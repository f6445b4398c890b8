"""Reference implementation names served by the fused GEMM+ReduceScatter kernel
(see ``ddlb_b200/primitives/TPColumnwise/compat.py`` for the policy).

* ``fuser`` (reference ``TPRowwise/fuser.py:171-312``): ``algorithm`` / ``s`` map
  one-to-one; the transport/stream options are accepted and recorded only.
* ``transformer_engine`` (reference ``TPRowwise/transformer_engine.py:66-79``):
  GEMM + reduce-scatter with sequence parallelism == ``default`` schedule.
"""
from __future__ import annotations

from .fused import FusedTPRowwise


class _CompatRowwise(FusedTPRowwise):
    COMPAT_DEFAULTS = {}
    COMPAT_ALLOWED = {}
    KERNEL_OPTIONS = FusedTPRowwise.DEFAULT_OPTIONS
    KERNEL_ALLOWED = FusedTPRowwise.ALLOWED_VALUES

    def __init__(self, m, n, k, dtype="float32", seed=42, **kwargs):
        from ..options import OptionsManager
        compat = OptionsManager(self.COMPAT_DEFAULTS, self.COMPAT_ALLOWED).parse(kwargs)
        self.compat_options = compat.as_dict()
        super().__init__(m, n, k, dtype=dtype, seed=seed, **self._translate(self.compat_options))

    def _translate(self, opts):
        return {}


class FuserTPRowwise(_CompatRowwise):
    COMPAT_DEFAULTS = {
        "backend": "nccl", "algorithm": "default", "s": 8,
        "offset_stream_indexing_by_rank": True, "multicast_protocol": "default",
        "inter_stream_synchronization": False,
    }
    COMPAT_ALLOWED = {
        "backend": ["nccl", "ucc", "ucc/tl/nccl", "ucc/tl/cuda", "cuda"],
        "algorithm": ["default", "coll_pipeline", "p2p_pipeline"],
        "s": (1, 1024),
        "offset_stream_indexing_by_rank": [True, False],
        "multicast_protocol": ["default", "memcpy", "multimem", "batch_memcpy"],
        "inter_stream_synchronization": [True, False],
    }
    DEFAULT_OPTIONS = dict(COMPAT_DEFAULTS)
    ALLOWED_VALUES = COMPAT_ALLOWED

    def _translate(self, o):
        return {"algorithm": o["algorithm"], "s": o["s"]}


class TransformerEngineTPRowwise(_CompatRowwise):
    DEFAULT_OPTIONS = {}

    def _translate(self, o):
        return {"algorithm": "default"}
