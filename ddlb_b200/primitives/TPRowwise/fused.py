"""``fused``: the product implementation of ``tp_rowwise`` — GEMM and ReduceScatter
in ONE hand-written sm_100a kernel per rank (``csrc/tp_gemm_sm100.cuh``).

Replaces cuBLAS + ``ncclReduceScatter`` (reference
``ddlb/primitives/TPRowwise/pytorch.py:70-85``) and the nvFuser pipelines
(``ddlb/primitives/TPRowwise/fuser.py:15-169``). Partial-output tiles leave TMEM, are
rounded once to the output type and stored with 128-bit stores into a staging slot; the
aux warps of every CTA of the owning rank sum the d slots in fp32 as soon as all of them
have landed. No full ``[m,n]`` partial is ever written to local HBM by a separate GEMM and
no separate reduction kernel runs.

``algorithm``: ``default`` and ``p2p_pipeline`` are the SAME schedule here (owners visited
ring-wise, one chunk per owner: that is what the reference's p2p pipeline does, and a
non-overlapped "GEMM, then collective" order has no place in a fused kernel);
``coll_pipeline`` splits every owner's rows into ``s`` slabs and visits slab j of all owners
before slab j+1.
``transport``: ``push`` (default/``auto``/``p2p``: partials go to the owner's staging, own
tiles last), ``pull`` (partials stay local, owners load them over NVLink), ``multimem``
(``multimem.ld_reduce``: the NVSwitch adds the d copies).
``tile``: ``default`` or ``wide`` (256 x 512 per CTA pair, see the tp_columnwise twin).
"""
from __future__ import annotations

from .tp_rowwise import TPRowwise


class FusedTPRowwise(TPRowwise):
    DEFAULT_OPTIONS = {
        "algorithm": "default",
        "s": 2,
        "cta_group": 0,
        "block_n": 0,
        "transport": "auto",
        "rank_rotate": True,
        "fp32_mode": "3xtf32",
        "tile": "default",
    }
    ALLOWED_VALUES = {
        "algorithm": ["default", "coll_pipeline", "p2p_pipeline"],
        "s": (1, 1024),
        "cta_group": [0, 1, 2],
        "block_n": [0, 128, 256],
        "transport": ["auto", "p2p", "push", "pull", "multimem"],
        "rank_rotate": [True, False],
        "fp32_mode": ["3xtf32", "tf32"],
        "tile": ["default", "wide"],
    }

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        import torch
        from ...ops.tp_gemm import FusedGemmReduceScatter

        comm = self.communicator
        if not comm.is_cuda:
            raise RuntimeError("the fused implementation needs a CUDA device (sm_100a)")
        if self.torch_dtype not in (torch.float16, torch.bfloat16, torch.float32):
            raise ValueError(
                f"fused tp_rowwise supports float16/bfloat16/float32 (got {self.dtype}); "
                "use the pytorch baseline for other dtypes")
        d = comm.world_size
        rows = self.m // d
        chunk_rows = 0
        if self.options["algorithm"] == "coll_pipeline":
            s = int(self.options["s"])
            if rows % s != 0:
                raise ValueError(f"rows per rank ({rows}) must be divisible by s ({s})")
            chunk_rows = rows // s
        heap = None
        if d > 1:
            heap = comm.symmetric_heap(2 * self.m * ((self.n + 255) // 256 * 256) * self.A.element_size() + (64 << 20))
        self.op = FusedGemmReduceScatter(
            heap, comm.rank, d, self.m, self.n, self.k, self.torch_dtype, self.A, self.B,
            chunk_rows=chunk_rows, cta_group=int(self.options["cta_group"]),
            block_n=int(self.options["block_n"]), transport=self.options["transport"],
            fp32_mode=self.options["fp32_mode"], unrotated=not self.options["rank_rotate"],
            wide_n=True if self.options["tile"] == "wide" else None)
        torch.cuda.synchronize()
        if d > 1:
            heap.barrier()
            torch.cuda.synchronize()

    def poison(self) -> None:
        op, comm = self.op, self.communicator
        op.out.fill_(float("nan"))
        if comm.world_size > 1:
            op.staging.fill_(float("nan"))               # both parities: nobody is mid-launch here
            comm.symmetric_heap().barrier()              # peers' pushes must not precede the fill

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
