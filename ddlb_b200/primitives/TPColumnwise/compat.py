"""Reference implementation NAMES served by the fused kernel.

Configs written for the reference name ``fuser`` (nvFuser), ``transformer_engine``
and ``jax`` implementations. None of those libraries is wrapped here; so that such
configs keep running, the names are accepted with the reference's option tables and
executed by the native fused kernel. A result row says so: its label reads
``fused[as fuser] (...)`` and options that have no equivalent here are listed under
``ignored:`` in the ``option`` column instead of decorating the label
(``ddlb_b200/benchmark.py:describe_implementation``).

* ``fuser`` (reference ``TPColumnwise/fuser.py:149-284``): ``algorithm`` / ``s`` / ``order``
  map one-to-one. ``offset_stream_indexing_by_rank=False`` selects the un-rotated shard
  order (every rank starts with shard 0). ``multicast_protocol`` selects the data path when
  ``backend=cuda`` (nvFuser's symmetric-memory backend): ``multimem`` -> ``multimem.st``
  through the NVSwitch, ``memcpy`` / ``batch_memcpy`` -> the push copies (every rank writes
  its shard into the peers' buffers, as a peer memcpy does), ``default`` -> the pull
  transport. ``backend`` values other than ``cuda`` (NCCL / UCC transports) and
  ``inter_stream_synchronization`` (one kernel has no streams to synchronise) have no
  equivalent and are reported as ignored.
* ``transformer_engine`` (reference ``TPColumnwise/transformer_engine.py:58-71``):
  userbuffers ring all-gather overlap == ``p2p_pipeline``.
* ``jax`` (reference ``TPColumnwise/jax_tp.py:60-76``): compiler-chosen all-gather
  + GEMM == ``default`` schedule.
"""
from __future__ import annotations

from .fused import FusedTPColumnwise


class _CompatColumnwise(FusedTPColumnwise):
    """Translate a reference option table into the fused kernel's options."""

    SERVED_BY = "fused"
    COMPAT_DEFAULTS = {}
    COMPAT_ALLOWED = {}
    KERNEL_OPTIONS = FusedTPColumnwise.DEFAULT_OPTIONS
    KERNEL_ALLOWED = FusedTPColumnwise.ALLOWED_VALUES

    def __init__(self, m, n, k, dtype="float32", seed=42, **kwargs):
        from ..options import OptionsManager
        compat = OptionsManager(self.COMPAT_DEFAULTS, self.COMPAT_ALLOWED).parse(kwargs)
        self.compat_options = compat.as_dict()
        super().__init__(m, n, k, dtype=dtype, seed=seed, **self._translate(self.compat_options))

    @classmethod
    def honoured_options(cls, options):
        """Names of the options that change what runs (the rest is reported as ignored)."""
        return set(options)

    @classmethod
    def _translate(cls, opts):
        return {}


def _fuser_transport(o, order_key="order"):
    """nvFuser's symmetric-memory protocols -> data path of the fused kernel."""
    if o.get("backend") != "cuda":
        return "auto"
    proto = o.get("multicast_protocol", "default")
    after = o.get(order_key) == "AG_after"
    if proto == "multimem":
        return "multimem"
    if proto in ("memcpy", "batch_memcpy"):
        return "st128" if after else "push"
    return "auto"


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

    @classmethod
    def honoured_options(cls, options):
        keep = {"order", "algorithm", "offset_stream_indexing_by_rank"}
        if options.get("algorithm") == "coll_pipeline":
            keep.add("s")
        if options.get("backend") == "cuda":
            keep |= {"backend", "multicast_protocol"}
        return keep

    @classmethod
    def _translate(cls, o):
        return {"order": o["order"], "algorithm": o["algorithm"], "s": o["s"],
                "rank_rotate": bool(o["offset_stream_indexing_by_rank"]),
                "transport": _fuser_transport(o)}


class TransformerEngineTPColumnwise(_CompatColumnwise):
    DEFAULT_OPTIONS = {}

    @classmethod
    def _translate(cls, o):
        return {"order": "AG_before", "algorithm": "p2p_pipeline"}


class JAXTPColumnwise(_CompatColumnwise):
    DEFAULT_OPTIONS = {}

    @classmethod
    def _translate(cls, o):
        return {"order": "AG_before", "algorithm": "default"}
