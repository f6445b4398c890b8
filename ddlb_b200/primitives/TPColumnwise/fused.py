"""``fused``: the product implementation of ``tp_columnwise`` — AllGather and GEMM
in ONE hand-written sm_100a kernel per rank (``csrc/tp_gemm_sm100.cuh``).

It replaces the library pair ``ncclAllGather`` + cuBLAS of the reference's
``pytorch`` implementation (``ddlb/primitives/TPColumnwise/pytorch.py:85-104``)
and the stream-pipelined nvFuser / Transformer-Engine variants
(``.../fuser.py:16-146``, ``.../transformer_engine.py:51-71``). Their knobs map to
tile-schedule parameters of the kernel:

=================  ==========================================================
``order``          ``AG_before``: the A shards cross NVLink into a local gathered
                   buffer while the GEMM consumes the chunks that have arrived.
                   ``AG_after``: each rank multiplies its shard; the epilogue writes
                   the C tiles into every rank's C.
``algorithm``      ``default``: ~1 MiB chunks, all peers in flight, stage-major.
                   ``coll_pipeline``: ``s`` chunks per shard, stage-major
                   (chunk j of every peer, then chunk j+1 ...).
                   ``p2p_pipeline``: one chunk per peer, peers visited in ring
                   order starting at rank+1 (d stages).
``s``              number of stages for ``coll_pipeline``.
``transport``      AG_before — ``pull`` (default, also ``auto``/``p2p``): the aux warps of
                   EVERY CTA read the peers' shards with 128-bit loads through registers;
                   ``push``: they write the own shard into every peer's buffer;
                   ``multimem``: one ``multimem.st`` per 16 bytes, replicated by the NVSwitch.
                   AG_after — ``tma`` (default/``auto``): TMA stores to all ranks; ``st128``:
                   128-bit stores to the peers; ``multimem``: one ``multimem.st``.
``rank_rotate``    ``False``: every rank visits the shards 0..d-1 in the same order
                   (nvFuser ``offset_stream_indexing_by_rank=False``).
``cta_group``      tcgen05 ``cta_group`` (0 = auto: 2 when rows/rank % 256 == 0).
``fp32_mode``      float32 only: ``3xtf32`` (default, ~fp32 accuracy) or ``tf32``.
``tile``           ``default`` (256 x 256 per CTA pair) or ``wide`` (256 x 512: two
                   accumulators share every A stage; fp16/bf16, cta_group 2, an even number
                   of 256-column blocks; measured 0.80 - 1.03x on one GPU, multi-rank runs
                   of it were checked in single-GPU emulation only).
=================  ==========================================================
"""
from __future__ import annotations

from .tp_columnwise import TPColumnwise


class FusedTPColumnwise(TPColumnwise):
    DEFAULT_OPTIONS = {
        "order": "AG_before",
        "algorithm": "default",
        "s": 8,
        "cta_group": 0,
        "block_n": 0,
        "transport": "auto",
        "rank_rotate": True,
        "fp32_mode": "3xtf32",
        "tile": "default",
    }
    ALLOWED_VALUES = {
        "order": ["AG_before", "AG_after"],
        "algorithm": ["default", "coll_pipeline", "p2p_pipeline"],
        "s": (1, 1024),
        "cta_group": [0, 1, 2],
        "block_n": [0, 128, 256],
        "transport": ["auto", "p2p", "pull", "push", "multimem", "tma", "st128"],
        "rank_rotate": [True, False],
        "fp32_mode": ["3xtf32", "tf32"],
        "tile": ["default", "wide"],
    }

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        import torch
        from ...ops.tp_gemm import FusedAllGatherGemm

        comm = self.communicator
        if not comm.is_cuda:
            raise RuntimeError("the fused implementation needs a CUDA device (sm_100a)")
        if self.torch_dtype not in (torch.float16, torch.bfloat16, torch.float32):
            raise ValueError(
                f"fused tp_columnwise supports float16/bfloat16/float32 (got {self.dtype}); "
                "use the pytorch baseline for other dtypes")
        d = comm.world_size
        rows = self.m // d
        chunk_rows = self._chunk_rows(rows)
        elt = 4 if self.torch_dtype == torch.float32 else 2
        need = self.m * max(self.k, 2 * self.n) * elt + (64 << 20)
        heap = comm.symmetric_heap(need) if d > 1 else None
        self.op = FusedAllGatherGemm(
            heap, comm.rank, d, self.m, self.n, self.k, self.torch_dtype, self.B,
            order=self.options["order"], chunk_rows=chunk_rows,
            cta_group=int(self.options["cta_group"]), block_n=int(self.options["block_n"]),
            transport=self.options["transport"], fp32_mode=self.options["fp32_mode"],
            unrotated=not self.options["rank_rotate"],
            wide_n=True if self.options["tile"] == "wide" else None,
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
            cta_group=int(self.options["cta_group"]), block_n=int(self.options["block_n"]),
            fp32_mode=self.options["fp32_mode"],
            wide_n=True if self.options["tile"] == "wide" else None)

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

    def poison(self) -> None:
        import torch
        op, comm = self.op, self.communicator
        if comm.world_size > 1:
            if op.a_gathered is not None:                # AG_before: the peers' rows
                keep = op.a_shard.clone()
                op.a_gathered.fill_(float("nan"))
                op.a_shard.copy_(keep)
            comm.symmetric_heap().barrier()              # peers' pushes must not precede the fill
        for out in op.out:
            out.fill_(float("nan"))
        if comm.world_size > 1:
            comm.symmetric_heap().barrier()

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
