"""``tp_columnwise``: AllGather -> GEMM.

``A`` is row (sequence) sharded ``[m/d, k]``, ``B`` ``[k, n]`` is replicated and
identical on every rank, the result ``[m, n]`` is replicated
(reference ``ddlb/primitives/TPColumnwise/tp_columnwise.py:13-22,99-124``).
"""
from __future__ import annotations

from ..base import TPPrimitive


class TPColumnwise(TPPrimitive):
    PRIMITIVE = "tp_columnwise"

    def _check_divisibility(self) -> None:
        d = self.communicator.world_size
        if self.m % d != 0:
            raise ValueError(
                f"Matrix dimension m ({self.m}) must be divisible by world_size ({d})")

    def _input_setup(self) -> None:
        import torch
        comm = self.communicator
        # same seed on every rank => identical unsharded tensors, no broadcast
        self.A_unsharded = 2 * torch.rand(self.m, self.k, dtype=self.torch_dtype, device="cpu") - 1
        rows = self.m // comm.world_size
        lo = comm.rank * rows
        self.A = self.A_unsharded[lo:lo + rows, :].to(comm.device).contiguous()
        self.B = 2 * torch.rand(self.k, self.n, dtype=self.torch_dtype, device=comm.device) - 1

    def reference_result(self):
        return self._oracle_matmul(self.A_unsharded, self.B)

    def comm_bytes_per_gpu(self) -> float:
        d = self.communicator.world_size
        return (d - 1) / d * self.m * self.k * self.A.element_size()
