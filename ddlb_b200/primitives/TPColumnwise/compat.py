"""Reference implementation names served by the fused kernel.

Configs written for the reference name ``fuser`` (nvFuser), ``transformer_engine``
and ``jax`` implementations. None of those libraries is wrapped here; so that such
configs keep running, the names are accepted with the reference's option tables and
executed by the native fused kernel with the overlap schedule the named back-end
would have used:

* ``fuser`` (reference ``TPColumnwise/fuser.py:149-284``): ``algorithm`` / ``s`` /
  ``order`` map one-to-one; ``backend``, ``multicast_protocol``,
  ``offset_stream_indexing_by_rank`` and ``inter_stream_synchronization`` describe
  nvFuser's stream/transport plumbing and are accepted and recorded only (peer
  order is always rank-rotated, transport is P2P over NVLink).
* ``transformer_engine`` (reference ``TPColumnwise/transformer_engine.py:58-71``):
  userbuffers ring all-gather overlap == ``p2p_pipeline``.
* ``jax`` (reference ``TPColumnwise/jax_tp.py:60-76``): compiler-chosen all-gather
  + GEMM == ``default`` schedule.
"""
from __future__ import annotations

from .fused import FusedTPColumnwise


class _CompatColumnwise(FusedTPColumnwise):
    """Translate a reference option table into the fused kernel's options."""

    COMPAT_DEFAULTS = {}
    COMPAT_ALLOWED = {}
    KERNEL_OPTIONS = FusedTPColumnwise.DEFAULT_OPTIONS
    KERNEL_ALLOWED = FusedTPColumnwise.ALLOWED_VALUES

    def __init__(self, m, n, k, dtype="float32", seed=42, **kwargs):
        from ..options import OptionsManager
        compat = OptionsManager(self.COMPAT_DEFAULTS, self.COMPAT_ALLOWED).parse(kwargs)
        self.compat_options = compat.as_dict()
        super().__init__(m, n, k, dtype=dtype, seed=seed, **self._translate(self.compat_options))

    def _translate(self, opts):
        return {}


class FuserTPColumnwise(_CompatColumnwise):
    COMPAT_DEFAULTS = {
        "backend": "nccl", "order": "AG_before", "algorithm": "default", "s": 8,
        "offset_stream_indexing_by_rank": True, "multicast_protocol": "default",
        "inter_stream_synchronization": False,
    }
    COMPAT_ALLOWED = {
        "backend": ["nccl", "ucc", "ucc/tl/nccl", "ucc/tl/cuda", "cuda"],
        "order": ["AG_before", "AG_after"],
        "algorithm": ["default", "coll_pipeline", "p2p_pipeline"],
        "s": (1, 1024),
        "offset_stream_indexing_by_rank": [True, False],
        "multicast_protocol": ["default", "memcpy", "multimem", "batch_memcpy"],
        "inter_stream_synchronization": [True, False],
    }
    DEFAULT_OPTIONS = dict(COMPAT_DEFAULTS)
    ALLOWED_VALUES = COMPAT_ALLOWED   # what the harness shows / filters on

    def _translate(self, o):
        return {"order": o["order"], "algorithm": o["algorithm"], "s": o["s"]}


class TransformerEngineTPColumnwise(_CompatColumnwise):
    DEFAULT_OPTIONS = {}

    def _translate(self, o):
        return {"order": "AG_before", "algorithm": "p2p_pipeline"}


class JAXTPColumnwise(_CompatColumnwise):
    DEFAULT_OPTIONS = {}

    def _translate(self, o):
        return {"order": "AG_before", "algorithm": "default"}
