"""Python front of the fused tcgen05 GEMM kernels (``csrc/tp_gemm_sm100.cuh``).

``TpGemmPlan`` wraps one ``ddlb_tp_gemm_plan``: tensor maps and launch shape are
resolved once, ``run()`` is a single kernel launch on the current stream.
``FusedAllGatherGemm`` / ``FusedGemmReduceScatter`` add the symmetric buffers and
signal words a multi-rank op needs; ``gemm()`` is the single-GPU convenience call.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence

from .. import _native as N

def _dtype_code(t) -> int:
    import torch
    if t == torch.float16:
        return N.DTYPE_F16
    if t == torch.bfloat16:
        return N.DTYPE_BF16
    if t == torch.float32:
        if not tf32_enabled():
            raise TypeError(
                "float32 runs through the kind::tf32 kernels, which have not been validated on "
                "hardware yet; set DDLB_ENABLE_TF32=1 to use them, or use the pytorch baseline")
        return N.DTYPE_F32
    raise TypeError(f"the tcgen05 kernels take float16/bfloat16 (and float32 via TF32), got {t}")


def tf32_enabled() -> bool:
    """float32 inputs via kind::tf32 MMAs: opt-in until validated on a B200."""
    import os
    return os.environ.get("DDLB_ENABLE_TF32", "0") == "1"


class TpGemmPlan:
    """One resolved kernel launch. Keeps references to the tensors it points at."""

    def __init__(self, cfg: N.TpGemmConfig, keepalive: Sequence = ()):
        self._lib = N.lib()
        self.cfg = cfg
        self._keep = list(keepalive)
        handle = ctypes.c_void_p()
        N.check(self._lib.ddlb_tp_gemm_create(ctypes.byref(cfg), ctypes.byref(handle)),
                "tp_gemm_create")
        self._h = handle
        info = (ctypes.c_int32 * 8)()
        N.check(self._lib.ddlb_tp_gemm_info(self._h, ctypes.byref(info)), "tp_gemm_info")
        self.info: Dict[str, int] = dict(zip(
            ("grid", "cluster", "smem_bytes", "cta_group", "block_n", "stages", "n_comm_ctas",
             "chunk_rows"), list(info)))
        self.epoch = 0
        self.launches = 0

    def run(self, stream=None) -> None:
        import torch
        s = stream if stream is not None else torch.cuda.current_stream()
        self.epoch += 1
        self.launches += 1
        N.check(self._lib.ddlb_tp_gemm_run(self._h, self.epoch, ctypes.c_void_p(s.cuda_stream)),
                "tp_gemm_run")

    def close(self) -> None:
        if self._h:
            self._lib.ddlb_tp_gemm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _base_config(mode: int, dtype, m: int, n: int, k: int, rank: int, world: int,
                 b_layout: int, cta_group: int, block_n: int, group_m: int, num_sms: int,
                 n_comm_ctas: int = 0, chunk_rows: int = 0, comm_transport: int = 0) -> N.TpGemmConfig:
    cfg = N.TpGemmConfig()
    cfg.mode, cfg.dtype, cfg.b_layout = mode, _dtype_code(dtype), b_layout
    cfg.cta_group, cfg.block_n, cfg.group_m, cfg.num_sms = cta_group, block_n, group_m, num_sms
    cfg.n_comm_ctas, cfg.rank, cfg.world, cfg.chunk_rows = n_comm_ctas, rank, world, chunk_rows
    cfg.comm_transport = comm_transport
    cfg.m, cfg.n, cfg.k = m, n, k
    return cfg


def plain_gemm_plan(a, b, out=None, *, b_is_nk: bool = False, cta_group: int = 0,
                    block_n: int = 0, group_m: int = 0, num_sms: int = 0):
    """Plan for ``out[m,n] = a[m,k] @ b`` with ``b`` ``[k,n]`` (or ``[n,k]`` if ``b_is_nk``)."""
    import torch
    assert a.is_cuda and b.is_cuda and a.is_contiguous() and b.is_contiguous()
    assert a.dtype == b.dtype
    m, k = a.shape
    n = b.shape[0] if b_is_nk else b.shape[1]
    assert (b.shape[1] if b_is_nk else b.shape[0]) == k
    if out is None:
        out = torch.empty(m, n, dtype=a.dtype, device=a.device)
    cfg = _base_config(N.MODE_PLAIN, a.dtype, m, n, k, 0, 1, N.B_NK if b_is_nk else N.B_KN,
                       cta_group, block_n, group_m, num_sms)
    cfg.a, cfg.b = a.data_ptr(), b.data_ptr()
    cfg.out[0] = out.data_ptr()
    return TpGemmPlan(cfg, keepalive=(a, b, out)), out


def gemm(a, b, out=None, **kw):
    """``a @ b`` on the tcgen05 kernel (single launch; plan is not cached)."""
    plan, out = plain_gemm_plan(a, b, out, **kw)
    plan.run()
    return out


class _FusedBase:
    """Symmetric buffers + signal words shared by the fused multi-rank ops."""

    def __init__(self, heap, rank: int, world: int):
        self.heap, self.rank, self.world = heap, rank, world
        self.plan: Optional[TpGemmPlan] = None

    def _alloc_signals(self, cfg: N.TpGemmConfig) -> None:
        words = N.lib().ddlb_tp_gemm_sig_words(ctypes.byref(cfg))
        if words < 0:
            N.check(1, "tp_gemm_sig_words")
        off = self.heap.alloc_signals(int(words))
        cfg.sig_local = self.heap.ptr(off)
        for p in range(self.world):
            cfg.sig_peer[p] = self.heap.ptr(off, p)

    @property
    def launches(self) -> int:
        return self.plan.launches if self.plan else 0

    def close(self) -> None:
        if self.plan is not None:
            self.plan.close()
            self.plan = None


class FusedAllGatherGemm(_FusedBase):
    """``C[m,n] = allgather_rows(A_shard)[m,k] @ B[k,n]`` in ONE kernel per rank.

    ``order="AG_before"``: the peers' shards are moved over NVLink into the local gathered
    buffer chunk by chunk (``transport``: by two warps of every CTA through registers, by a
    TMA warp per CTA, or by dedicated comm CTAs) while the GEMM multiplies the chunks that
    have arrived. ``order="AG_after"``: each rank multiplies its own shard and its epilogue
    stores the C tiles into every rank's output.
    The rank's input shard must be written into ``self.a_shard`` (a view of the
    symmetric gathered buffer).
    """

    def __init__(self, heap, rank: int, world: int, m: int, n: int, k: int, dtype, b,
                 order: str = "AG_before", chunk_rows: int = 0, n_comm_ctas: int = 0,
                 cta_group: int = 0, block_n: int = 0, group_m: int = 0, num_sms: int = 0,
                 b_is_nk: bool = False, transport: str = "cta_pull", trace: bool = False,
                 variant: int = 0):
        super().__init__(heap, rank, world)
        import torch
        if m % world:
            raise ValueError(f"m ({m}) must be divisible by world ({world})")
        transports = {"cta": N.COMM_EVERY_CTA, "cta_tma": N.COMM_EVERY_CTA_TMA,
                      "cta_pull": N.COMM_EVERY_CTA_PULL,
                      "pull": N.COMM_PULL_CTAS, "push": N.COMM_PUSH_CTAS}
        if transport not in transports:
            raise ValueError(f"transport must be one of {sorted(transports)}, got {transport!r}")
        rows = m // world
        self.order = order
        layout = N.B_NK if b_is_nk else N.B_KN
        if order == "AG_before":
            self.out = [torch.empty(m, n, dtype=dtype, device=b.device)]
            cfg = _base_config(N.MODE_AG_BEFORE if world > 1 else N.MODE_PLAIN, dtype, m, n, k,
                               rank, world, layout, cta_group, block_n, group_m, num_sms,
                               n_comm_ctas, chunk_rows, transports[transport])
            if world == 1:
                # nothing to gather: an ordinary device buffer, no symmetric heap needed
                self.a_gathered = torch.empty(m, k, dtype=dtype, device=b.device)
                cfg.a = self.a_gathered.data_ptr()
            else:
                a_off, self.a_gathered = heap.empty((m, k), dtype)
                cfg.a = heap.ptr(a_off)
                for p in range(world):
                    cfg.peer_a[p] = heap.ptr(a_off, p)
            self.a_shard = self.a_gathered[rank * rows:(rank + 1) * rows]
            cfg.out[rank] = self.out[0].data_ptr()
        elif order == "AG_after":
            self.a_shard = torch.empty(rows, k, dtype=dtype, device=b.device)
            self.a_gathered = None
            if world > 1:
                # outputs are double-buffered by launch parity: a peer may run one launch ahead
                offs = [heap.empty((m, n), dtype) for _ in range(2)]
                self.out = [t for _, t in offs]
                cfgs = []
                for off, _ in offs:
                    cfg = _base_config(N.MODE_AG_AFTER, dtype, m, n, k, rank, world, layout,
                                       cta_group, block_n, group_m, num_sms)
                    cfg.a = self.a_shard.data_ptr()
                    for p in range(world):
                        cfg.out[p] = heap.ptr(off, p)
                    cfgs.append(cfg)
                cfg = cfgs
            else:
                self.out = [torch.empty(m, n, dtype=dtype, device=b.device)]
                cfg = _base_config(N.MODE_PLAIN, dtype, m, n, k, 0, 1, layout, cta_group, block_n,
                                   group_m, num_sms)
                cfg.a = self.a_shard.data_ptr()
                cfg.out[0] = self.out[0].data_ptr()
        else:
            raise ValueError(f"unknown order {order!r}")
        self.b = b
        cfgs = cfg if isinstance(cfg, list) else [cfg]
        self.trace_buf = None
        if trace:
            from ..utils.trace import alloc_trace_buffer
            self.trace_buf = alloc_trace_buffer(b.device)
        for c in cfgs:
            c.b = b.data_ptr()
            c.variant = int(variant)          # experimental A/B schedule variants (0 = production)
            if self.trace_buf is not None:
                c.trace = self.trace_buf.data_ptr()
        if world > 1:
            self._alloc_signals(cfgs[0])
            for c in cfgs[1:]:
                c.sig_local = cfgs[0].sig_local
                for p in range(world):
                    c.sig_peer[p] = cfgs[0].sig_peer[p]
        self.plans = [TpGemmPlan(c, keepalive=(self.a_shard, b, *self.out)) for c in cfgs]
        self.plan = self.plans[0]
        self._epoch = 0

    def run(self, stream=None):
        import torch
        s = stream if stream is not None else torch.cuda.current_stream()
        self._epoch += 1
        idx = self._epoch % len(self.plans)
        plan = self.plans[idx]
        plan.epoch = self._epoch - 1          # plans share one signal block: one epoch counter
        plan.run(s)
        return self.out[idx]

    @property
    def launches(self) -> int:
        return sum(p.launches for p in self.plans)

    def close(self) -> None:
        for p in self.plans:
            p.close()
        self.plans = []
        self.plan = None


class FusedGemmReduceScatter(_FusedBase):
    """``C_shard[m/d,n] = reduce_scatter_rows(A[m,k/d] @ B[k/d,n])`` in ONE kernel per rank.

    Tiles are visited owner by owner starting with the next rank; a tile that belongs
    to another rank is stored (TMA, P2P) into that rank's staging buffer and its
    arrival counter bumped; the rank's own tiles come last and their epilogue adds
    the d-1 staged partials in fp32 before the single rounding to the output type.
    """

    def __init__(self, heap, rank: int, world: int, m: int, n: int, k: int, dtype, a, b,
                 chunk_rows: int = 0, cta_group: int = 0, block_n: int = 0, group_m: int = 0,
                 num_sms: int = 0, b_is_nk: bool = False, trace: bool = False, variant: int = 0):
        super().__init__(heap, rank, world)
        import torch
        self.trace_buf = None
        if trace:
            from ..utils.trace import alloc_trace_buffer
            self.trace_buf = alloc_trace_buffer(a.device)
        if m % world or k % world:
            raise ValueError(f"m ({m}) and k ({k}) must be divisible by world ({world})")
        rows = m // world
        self.a, self.b = a, b
        self.out = torch.empty(rows, n, dtype=dtype, device=a.device)
        layout = N.B_NK if b_is_nk else N.B_KN
        if world == 1:
            cfg = _base_config(N.MODE_PLAIN, dtype, m, n, k, 0, 1, layout, cta_group, block_n,
                               group_m, num_sms)
            cfg.out[0] = self.out.data_ptr()
        else:
            st_off, self.staging = heap.empty((2, world, rows, n), dtype)
            cfg = _base_config(N.MODE_RS, dtype, m, n, k, rank, world, layout, cta_group, block_n,
                               group_m, num_sms, 0, chunk_rows)
            for p in range(world):
                cfg.out[p] = heap.ptr(st_off, p)
            cfg.out[rank] = self.out.data_ptr()
            cfg.stage_local = heap.ptr(st_off)
            self._alloc_signals(cfg)
        cfg.a, cfg.b = a.data_ptr(), b.data_ptr()
        cfg.variant = int(variant)            # experimental A/B schedule variants (0 = production)
        if self.trace_buf is not None:
            cfg.trace = self.trace_buf.data_ptr()
        self.plan = TpGemmPlan(cfg, keepalive=(a, b, self.out))

    def run(self, stream=None):
        self.plan.run(stream)
        return self.out
