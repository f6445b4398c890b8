"""Reference implementation NAMES served by the fused GEMM+ReduceScatter kernel
(see ``ddlb_b200/primitives/TPColumnwise/compat.py`` for the policy and the row labels).

* ``fuser`` (reference ``TPRowwise/fuser.py:171-312``): ``algorithm`` / ``s`` map
  one-to-one (``default`` and ``p2p_pipeline`` are the same ring schedule here, see
  ``fused.py``); ``offset_stream_indexing_by_rank=False`` selects the un-rotated owner order;
  with ``backend=cuda``, ``multicast_protocol=multimem`` selects ``multimem.ld_reduce`` and
  ``memcpy`` / ``batch_memcpy`` the push transport. Other ``backend`` values and
  ``inter_stream_synchronization`` are reported as ignored.
* ``transformer_engine`` (reference ``TPRowwise/transformer_engine.py:66-79``):
  GEMM + reduce-scatter with sequence parallelism == ``default`` schedule.
"""
from __future__ import annotations

from .fused import FusedTPRowwise


class _CompatRowwise(FusedTPRowwise):
    SERVED_BY = "fused"
    COMPAT_DEFAULTS = {}
    COMPAT_ALLOWED = {}
    KERNEL_OPTIONS = FusedTPRowwise.DEFAULT_OPTIONS
    KERNEL_ALLOWED = FusedTPRowwise.ALLOWED_VALUES

    def __init__(self, m, n, k, dtype="float32", seed=42, **kwargs):
        from ..options import OptionsManager
        compat = OptionsManager(self.COMPAT_DEFAULTS, self.COMPAT_ALLOWED).parse(kwargs)
        self.compat_options = compat.as_dict()
        super().__init__(m, n, k, dtype=dtype, seed=seed, **self._translate(self.compat_options))

    @classmethod
    def honoured_options(cls, options):
        return set(options)

    @classmethod
    def _translate(cls, opts):
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

    @classmethod
    def honoured_options(cls, options):
        keep = {"algorithm", "offset_stream_indexing_by_rank"}
        if options.get("algorithm") == "coll_pipeline":
            keep.add("s")
        if options.get("backend") == "cuda":
            keep |= {"backend", "multicast_protocol"}
        return keep

    @classmethod
    def _translate(cls, o):
        transport = "auto"
        if o["backend"] == "cuda":
            transport = {"multimem": "multimem", "memcpy": "push",
                         "batch_memcpy": "push"}.get(o["multicast_protocol"], "auto")
        return {"algorithm": o["algorithm"], "s": o["s"],
                "rank_rotate": bool(o["offset_stream_indexing_by_rank"]), "transport": transport}


class TransformerEngineTPRowwise(_CompatRowwise):
    DEFAULT_OPTIONS = {}

    @classmethod
    def _translate(cls, o):
        return {"algorithm": "default"}
