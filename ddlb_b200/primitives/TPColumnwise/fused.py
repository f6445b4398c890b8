"""``fused``: the product implementation of ``tp_columnwise`` — AllGather and GEMM
in ONE hand-written sm_100a kernel per rank (``csrc/tp_gemm_sm100.cuh``).

It replaces the library pair ``ncclAllGather`` + cuBLAS of the reference's
``pytorch`` implementation (``ddlb/primitives/TPColumnwise/pytorch.py:85-104``)
and the stream-pipelined nvFuser / Transformer-Engine variants
(``.../fuser.py:16-146``, ``.../transformer_engine.py:51-71``). Their knobs map to
tile-schedule parameters of the kernel:

=================  ==========================================================
``order``          ``AG_before``: peers' A shards are pulled over NVLink by comm
                   CTAs into a local gathered buffer while GEMM CTAs consume the
                   chunks that have arrived. ``AG_after``: each rank multiplies
                   its shard; the epilogue stores C tiles into every rank's C.
``algorithm``      ``default``: ~1 MiB chunks, all peers in flight, stage-major.
                   ``coll_pipeline``: ``s`` chunks per shard, stage-major
                   (chunk j of every peer, then chunk j+1 ...).
                   ``p2p_pipeline``: one chunk per peer, peers visited in ring
                   order starting at rank+1 (d stages).
``s``              number of stages for ``coll_pipeline``.
``comm_sms``       CTAs used for communication (0 = auto).
``cta_group``      tcgen05 ``cta_group`` (0 = auto: 2 when rows/rank % 256 == 0).
``transport``      ``cta_pull`` (default, also ``auto``/``p2p``): two warps of EVERY CTA pull the
                   peers' shards with 128-bit loads/stores through registers, so NVLink
                   traffic is spread over all SMs and the GEMM keeps every SM.
                   ``cta``: same warps, pushing the own shard to all peers. ``cta_tma``: one
                   warp of every CTA pulls 8 KiB pieces with bulk-async copies. ``pull`` /
                   ``push``: ``comm_sms`` dedicated CTAs (bulk-async copies).
=================  ==========================================================
"""
from __future__ import annotations

from .tp_columnwise import TPColumnwise


class FusedTPColumnwise(TPColumnwise):
    DEFAULT_OPTIONS = {
        "order": "AG_before",
        "algorithm": "default",
        "s": 8,
        "comm_sms": 0,
        "cta_group": 0,
        "block_n": 0,
        "transport": "auto",
    }
    ALLOWED_VALUES = {
        "order": ["AG_before", "AG_after"],
        "algorithm": ["default", "coll_pipeline", "p2p_pipeline"],
        "s": (1, 1024),
        "comm_sms": (0, 64),
        "cta_group": [0, 1, 2],
        "block_n": [0, 128, 256],
        "transport": ["auto", "cta", "cta_pull", "cta_tma", "push", "pull", "p2p"],
    }

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        import torch
        from ...ops.tp_gemm import FusedAllGatherGemm

        comm = self.communicator
        if not comm.is_cuda:
            raise RuntimeError("the fused implementation needs a CUDA device (sm_100a)")
        from ...ops.tp_gemm import tf32_enabled
        ok_dtypes = (torch.float16, torch.bfloat16) + ((torch.float32,) if tf32_enabled() else ())
        if self.torch_dtype not in ok_dtypes:
            raise ValueError(
                f"fused tp_columnwise supports float16/bfloat16 (got {self.dtype}); float32 through "
                "kind::tf32 is opt-in (DDLB_ENABLE_TF32=1) until validated on hardware; "
                "use the pytorch baseline for other dtypes")
        d = comm.world_size
        rows = self.m // d
        chunk_rows = self._chunk_rows(rows)
        elt = 4 if self.torch_dtype == torch.float32 else 2
        need = self.m * max(self.k, 2 * self.n) * elt + (64 << 20)
        heap = None
        if d > 1:
            comm.init_process_group()
            heap = comm.symmetric_heap(need)
        self.op = FusedAllGatherGemm(
            heap, comm.rank, d, self.m, self.n, self.k, self.torch_dtype, self.B,
            order=self.options["order"], chunk_rows=chunk_rows,
            n_comm_ctas=int(self.options["comm_sms"]), cta_group=int(self.options["cta_group"]),
            block_n=int(self.options["block_n"]),
            transport={"auto": "cta_pull", "p2p": "cta_pull"}.get(self.options["transport"],
                                                        self.options["transport"]),
        ) if d > 1 else self._single_rank_op()
        self.op.a_shard.copy_(self.A)
        self.A = self.op.a_shard
        torch.cuda.synchronize()
        if d > 1:
            heap.barrier()
            torch.cuda.synchronize()

    def _single_rank_op(self):
        from ...ops.tp_gemm import FusedAllGatherGemm
        return FusedAllGatherGemm(
            None, 0, 1, self.m, self.n, self.k, self.torch_dtype, self.B, order="AG_after",
            cta_group=int(self.options["cta_group"]), block_n=int(self.options["block_n"]))

    def _chunk_rows(self, rows: int) -> int:
        algo = self.options["algorithm"]
        if algo == "p2p_pipeline":
            return rows
        if algo == "coll_pipeline":
            s = int(self.options["s"])
            if rows % s != 0:
                raise ValueError(f"rows per rank ({rows}) must be divisible by s ({s})")
            return rows // s
        return 0

    def run(self):
        return self.op.run()

    def device_barrier(self) -> None:
        if self.communicator.world_size > 1:
            self.communicator.symmetric_heap().barrier()

    @property
    def gpu_launches(self) -> int:
        return self.op.launches

    def close(self) -> None:
        self.op.close()
