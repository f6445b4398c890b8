"""Python front of the fused tcgen05 GEMM kernels (``csrc/tp_gemm_sm100.cuh``).

``TpGemmPlan`` wraps one ``ddlb_tp_gemm_plan``: tensor maps and launch shape are
resolved once, ``run()`` is a single kernel launch on the current stream.
``FusedAllGatherGemm`` / ``FusedGemmReduceScatter`` add the symmetric buffers and
signal words a multi-rank op needs; ``gemm()`` is the single-GPU convenience call.

float32 runs on the tensor cores as 3xTF32 (``fp32_mode="3xtf32"``, default: both operands
are split into tf32 hi + lo parts in shared memory and three ``kind::tf32`` MMAs are issued
per k-step) or as one TF32 pass (``"tf32"``). Either way B must be ``[n,k]``; the wrappers
transpose a ``[k,n]`` weight ONCE at construction (the reference's Transformer-Engine
implementation does the same with its weight, ``TPColumnwise/transformer_engine.py:69-71``).
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Sequence

from .. import _native as N

AG_TRANSPORTS = {"pull": N.AG_PULL, "push": N.AG_PUSH, "multimem": N.AG_PUSH_MULTICAST}
EPI_TRANSPORTS = {"tma": N.EPI_TMA, "st128": N.EPI_ST128, "multimem": N.EPI_MULTICAST}
RS_TRANSPORTS = {"push": N.RS_PUSH, "pull": N.RS_PULL, "multimem": N.RS_PULL_MULTICAST}
FP32_MODES = {"3xtf32": N.FP32_3XTF32, "tf32": N.FP32_TF32}
ACTIVATIONS = {None: N.ACT_NONE, "none": N.ACT_NONE, "relu": N.ACT_RELU, "silu": N.ACT_SILU,
               "gelu_tanh": N.ACT_GELU_TANH}


def _dtype_code(t) -> int:
    import torch
    if t == torch.float16:
        return N.DTYPE_F16
    if t == torch.bfloat16:
        return N.DTYPE_BF16
    if t == torch.float32:
        return N.DTYPE_F32
    raise TypeError(f"the tcgen05 kernels take float16/bfloat16/float32, got {t}")


def _weight_for_kernel(b, b_is_nk: bool):
    """fp32 operands must be K-major for the tensor cores: ``[k,n]`` weights are transposed once."""
    import torch
    if b.dtype == torch.float32 and not b_is_nk:
        return b.t().contiguous(), True
    return b, b_is_nk


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
        info = (ctypes.c_int32 * 12)()
        N.check(self._lib.ddlb_tp_gemm_info(self._h, ctypes.byref(info)), "tp_gemm_info")
        self.info: Dict[str, int] = dict(zip(
            ("grid", "cluster", "smem_bytes", "cta_group", "block_n", "stages", "threads",
             "chunk_rows", "pairs", "wide", "pdl"), list(info)))
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
                 chunk_rows: int = 0, fp32_mode: str = "3xtf32",
                 unrotated: bool = False, activation: Optional[str] = None,
                 cluster_pairs: int = 0, pdl: Optional[bool] = None,
                 wide_n: Optional[bool] = None) -> N.TpGemmConfig:
    if fp32_mode not in FP32_MODES:
        raise ValueError(f"fp32_mode must be one of {sorted(FP32_MODES)}, got {fp32_mode!r}")
    if activation not in ACTIVATIONS:
        raise ValueError(f"activation must be one of {list(ACTIVATIONS)}, got {activation!r}")
    cfg = N.TpGemmConfig()
    cfg.mode, cfg.dtype, cfg.b_layout = mode, _dtype_code(dtype), b_layout
    cfg.cta_group, cfg.block_n, cfg.group_m, cfg.num_sms = cta_group, block_n, group_m, num_sms
    cfg.rank, cfg.world, cfg.chunk_rows = rank, world, chunk_rows
    cfg.fp32_mode = FP32_MODES[fp32_mode]
    cfg.unrotated = 1 if unrotated else 0
    cfg.epilogue_act = ACTIVATIONS[activation]
    cfg.cluster_pairs = int(cluster_pairs)
    cfg.pdl = 0 if pdl is None else (1 if pdl else 2)   # None: on for single-rank plans
    cfg.wide_n = 0 if wide_n is None else (1 if wide_n else 2)   # opt-in: auto never picks it
    cfg.m, cfg.n, cfg.k = m, n, k
    return cfg


def plain_gemm_plan(a, b, out=None, *, b_is_nk: bool = False, cta_group: int = 0,
                    block_n: int = 0, group_m: int = 0, num_sms: int = 0,
                    fp32_mode: str = "3xtf32", activation: Optional[str] = None,
                    cluster_pairs: int = 0, pdl: Optional[bool] = None,
                    wide_n: Optional[bool] = None):
    """Plan for ``out[m,n] = a[m,k] @ b`` with ``b`` ``[k,n]`` (or ``[n,k]`` if ``b_is_nk``).

    ``pdl`` (default on): the launch uses programmatic stream serialization, so the kernel's
    prologue overlaps the tail of the previous kernel of the stream; all global accesses still
    wait for that kernel to complete."""
    import torch
    assert a.is_cuda and b.is_cuda and a.is_contiguous() and b.is_contiguous()
    assert a.dtype == b.dtype
    m, k = a.shape
    n = b.shape[0] if b_is_nk else b.shape[1]
    assert (b.shape[1] if b_is_nk else b.shape[0]) == k
    if out is None:
        out = torch.empty(m, n, dtype=a.dtype, device=a.device)
    b, b_is_nk = _weight_for_kernel(b, b_is_nk)
    cfg = _base_config(N.MODE_PLAIN, a.dtype, m, n, k, 0, 1, N.B_NK if b_is_nk else N.B_KN,
                       cta_group, block_n, group_m, num_sms, fp32_mode=fp32_mode,
                       activation=activation, cluster_pairs=cluster_pairs, pdl=pdl, wide_n=wide_n)
    cfg.a, cfg.b = a.data_ptr(), b.data_ptr()
    cfg.out[0] = out.data_ptr()
    return TpGemmPlan(cfg, keepalive=(a, b, out)), out


def gemm(a, b, out=None, **kw):
    """``a @ b`` on the tcgen05 kernel (single launch; plan is not cached)."""
    plan, out = plain_gemm_plan(a, b, out, **kw)
    plan.run()
    return out


def _resolve_transport(name: str, table: dict, default: str, heap, what: str) -> str:
    """``auto`` -> the default; ``multimem`` needs a heap with an NVLS multicast window."""
    if name in ("auto", "p2p"):
        name = default
    if name not in table:
        raise ValueError(f"{what} transport must be one of {sorted(table)} (or auto), got {name!r}")
    if name == "multimem" and not getattr(heap, "has_multicast", False):
        raise RuntimeError(
            f"{what} transport 'multimem' needs the VMM symmetric heap with an NVLS multicast "
            "window (DDLB_SYMM_MODE=vmm or auto on NVSwitch systems); this heap has none")
    return name


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

    ``order="AG_before"``: the shards are moved over NVLink chunk by chunk by the aux warps
    of every CTA (``transport``: ``pull`` the peers' shards - default -, ``push`` the own
    shard to every peer, or ``multimem``: one ``multimem.st`` per 16 bytes replicated by the
    NVSwitch) while the GEMM multiplies the chunks that have arrived. The kernel returns only
    after every peer has finished reading this rank's shard, so the caller may overwrite
    ``a_shard`` right after it (the guarantee ``ncclAllGather`` gives for its send buffer).
    ``order="AG_after"``: each rank multiplies its own shard and its epilogue writes the C
    tiles into every rank's output (``transport``: ``tma`` - default -, ``st128``,
    ``multimem``).
    The rank's input shard must be written into ``self.a_shard`` (a view of the
    symmetric gathered buffer).
    """

    def __init__(self, heap, rank: int, world: int, m: int, n: int, k: int, dtype, b,
                 order: str = "AG_before", chunk_rows: int = 0,
                 cta_group: int = 0, block_n: int = 0, group_m: int = 0, num_sms: int = 0,
                 b_is_nk: bool = False, transport: str = "auto", trace: bool = False,
                 fp32_mode: str = "3xtf32", unrotated: bool = False,
                 activation: Optional[str] = None, out=None, comm_warps: int = 0,
                 cluster_pairs: int = 0, wide_n: Optional[bool] = None):
        super().__init__(heap, rank, world)
        import torch
        if m % world:
            raise ValueError(f"m ({m}) must be divisible by world ({world})")
        rows = m // world
        self.order = order
        b, b_is_nk = _weight_for_kernel(b, b_is_nk)
        layout = N.B_NK if b_is_nk else N.B_KN
        common = dict(fp32_mode=fp32_mode, unrotated=unrotated, activation=activation,
                      cluster_pairs=cluster_pairs, wide_n=wide_n)
        if out is not None and order != "AG_before":
            raise ValueError("a caller-provided output is supported for order='AG_before' only")
        if order == "AG_before":
            self.transport = _resolve_transport(transport, AG_TRANSPORTS, "pull", heap, "AG_before")
            self.out = [out if out is not None else torch.empty(m, n, dtype=dtype, device=b.device)]
            cfg = _base_config(N.MODE_AG_BEFORE if world > 1 else N.MODE_PLAIN, dtype, m, n, k,
                               rank, world, layout, cta_group, block_n, group_m, num_sms,
                               chunk_rows, **common)
            cfg.ag_transport = AG_TRANSPORTS[self.transport]
            cfg.comm_warps = int(comm_warps)
            if world == 1:
                # nothing to gather: an ordinary device buffer, no symmetric heap needed
                self.a_gathered = torch.empty(m, k, dtype=dtype, device=b.device)
                cfg.a = self.a_gathered.data_ptr()
            else:
                a_off, self.a_gathered = heap.empty((m, k), dtype)
                cfg.a = heap.ptr(a_off)
                for p in range(world):
                    cfg.peer_a[p] = heap.ptr(a_off, p)
                if self.transport == "multimem":
                    cfg.mc_ptr = heap.mc_ptr(a_off)
            self.a_shard = self.a_gathered[rank * rows:(rank + 1) * rows]
            cfg.out[rank] = self.out[0].data_ptr()
        elif order == "AG_after":
            self.transport = _resolve_transport(transport, EPI_TRANSPORTS, "tma", heap, "AG_after")
            self.a_shard = torch.empty(rows, k, dtype=dtype, device=b.device)
            self.a_gathered = None
            if world > 1:
                # outputs are double-buffered by launch parity: a peer may run one launch ahead
                offs = [heap.empty((m, n), dtype) for _ in range(2)]
                self.out = [t for _, t in offs]
                cfgs = []
                for off, _ in offs:
                    cfg = _base_config(N.MODE_AG_AFTER, dtype, m, n, k, rank, world, layout,
                                       cta_group, block_n, group_m, num_sms, **common)
                    cfg.epi_transport = EPI_TRANSPORTS[self.transport]
                    cfg.a = self.a_shard.data_ptr()
                    for p in range(world):
                        cfg.out[p] = heap.ptr(off, p)
                    if self.transport == "multimem":
                        cfg.mc_ptr = heap.mc_ptr(off)
                    cfgs.append(cfg)
                cfg = cfgs
            else:
                self.out = [torch.empty(m, n, dtype=dtype, device=b.device)]
                cfg = _base_config(N.MODE_PLAIN, dtype, m, n, k, 0, 1, layout, cta_group, block_n,
                                   group_m, num_sms, **common)
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

    Every partial tile leaves TMEM, is rounded once to the output type and stored with 128-bit
    stores into a staging slot; the owner's aux warps (all CTAs, work-queue) sum the d slots
    of each 32-row x 128-byte unit in fp32 as soon as all of them have landed and write the
    output shard. ``transport``:

    * ``push`` (default): slots live at the OWNER; owners are visited ring-wise starting with
      the next rank, the own tiles last (their partials need no NVLink); the reducers read
      local memory.
    * ``pull``: slots stay LOCAL; every rank visits the owners 0..d-1 in the same order and
      the owner's reducers load the d copies over NVLink.
    * ``multimem``: like ``pull``, but the owner issues ONE ``multimem.ld_reduce`` per
      16 bytes and the NVSwitch adds the d copies (needs the VMM heap's multicast window).
    """

    def __init__(self, heap, rank: int, world: int, m: int, n: int, k: int, dtype, a, b,
                 chunk_rows: int = 0, cta_group: int = 0, block_n: int = 0, group_m: int = 0,
                 num_sms: int = 0, b_is_nk: bool = False, trace: bool = False,
                 transport: str = "auto", fp32_mode: str = "3xtf32", unrotated: bool = False,
                 cluster_pairs: int = 0, wide_n: Optional[bool] = None):
        super().__init__(heap, rank, world)
        import torch
        self.trace_buf = None
        if trace:
            from ..utils.trace import alloc_trace_buffer
            self.trace_buf = alloc_trace_buffer(a.device)
        if m % world or k % world:
            raise ValueError(f"m ({m}) and k ({k}) must be divisible by world ({world})")
        rows = m // world
        b, b_is_nk = _weight_for_kernel(b, b_is_nk)
        self.a, self.b = a, b
        self.out = torch.empty(rows, n, dtype=dtype, device=a.device)
        layout = N.B_NK if b_is_nk else N.B_KN
        # auto: the in-switch reduction wherever the heap has an NVLS window (measured on B200,
        # headline shape, back to back: 71.7 / 55.6 / 79.4 us at 2 / 4 / 8 ranks against
        # 75.1 / 67.8 / 84.9 us for push), otherwise push
        default = "multimem" if getattr(heap, "has_multicast", False) else "push"
        self.transport = _resolve_transport(transport, RS_TRANSPORTS, default, heap, "tp_rowwise")
        if world == 1:
            cfg = _base_config(N.MODE_PLAIN, dtype, m, n, k, 0, 1, layout, cta_group, block_n,
                               group_m, num_sms, fp32_mode=fp32_mode, cluster_pairs=cluster_pairs,
                               wide_n=wide_n)
            cfg.out[0] = self.out.data_ptr()
        else:
            # staging is unit-major and covers whole 256-column blocks (see rs_reduce_loop)
            n_pad = (n + 255) // 256 * 256
            st_off, self.staging = heap.empty((2, world, rows, n_pad), dtype)
            cfg = _base_config(N.MODE_RS, dtype, m, n, k, rank, world, layout, cta_group, block_n,
                               group_m, num_sms, chunk_rows, fp32_mode=fp32_mode,
                               unrotated=unrotated, cluster_pairs=cluster_pairs, wide_n=wide_n)
            cfg.rs_transport = RS_TRANSPORTS[self.transport]
            for p in range(world):
                cfg.stage[p] = heap.ptr(st_off, p)
            if self.transport == "multimem":
                cfg.mc_ptr = heap.mc_ptr(st_off)
            cfg.out[rank] = self.out.data_ptr()
            self._alloc_signals(cfg)
        cfg.a, cfg.b = a.data_ptr(), b.data_ptr()
        if self.trace_buf is not None:
            cfg.trace = self.trace_buf.data_ptr()
        self.plan = TpGemmPlan(cfg, keepalive=(a, b, self.out))

    def run(self, stream=None):
        self.plan.run(stream)
        return self.out
