"""``tp_rowwise``: GEMM -> ReduceScatter (sequence-parallel output).

``A`` ``[m, k/d]`` and ``B`` ``[k/d, n]`` are sharded on the contracting
dimension; rank ``r`` returns rows ``[r*m/d, (r+1)*m/d)`` of ``A @ B``
(reference ``ddlb/primitives/TPRowwise/tp_rowwise.py:13-27,112-140,153-184``).
"""
from __future__ import annotations

from ..base import TPPrimitive


class TPRowwise(TPPrimitive):
    PRIMITIVE = "tp_rowwise"

    def _check_divisibility(self) -> None:
        d = self.communicator.world_size
        if self.k % d != 0:
            raise ValueError(
                f"Matrix dimension k ({self.k}) must be divisible by world_size ({d})")
        if self.m % d != 0:
            raise ValueError(
                f"Matrix dimension m ({self.m}) must be divisible by world_size ({d})")

    def _input_setup(self) -> None:
        import torch
        comm = self.communicator
        d, r = comm.world_size, comm.rank
        kk = self.k // d
        self.A_unsharded = 2 * torch.rand(self.m, self.k, dtype=self.torch_dtype, device="cpu") - 1
        self.A = self.A_unsharded[:, r * kk:(r + 1) * kk].to(comm.device).contiguous()
        self.B_unsharded = 2 * torch.rand(self.k, self.n, dtype=self.torch_dtype, device="cpu") - 1
        self.B = self.B_unsharded[r * kk:(r + 1) * kk, :].to(comm.device).contiguous()

    def reference_result(self):
        d, r = self.communicator.world_size, self.communicator.rank
        rows = self.m // d
        # only this rank's rows of the unsharded product are needed
        return self._oracle_matmul(self.A_unsharded[r * rows:(r + 1) * rows, :], self.B_unsharded)

    def comm_bytes_per_gpu(self) -> float:
        d = self.communicator.world_size
        return (d - 1) / d * self.m * self.n * self.A.element_size()
