"""Compute-only ceiling for ``tp_columnwise``: a cuBLAS matmul, no communication
(reference ``ddlb/primitives/TPColumnwise/compute_only.py:8-55``)."""
from __future__ import annotations

from .tp_columnwise import TPColumnwise


class ComputeOnlyTPColumnwise(TPColumnwise):
    DEFAULT_OPTIONS = {"size": "sharded"}
    ALLOWED_VALUES = {"size": ["sharded", "unsharded"]}

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.size = self.options["size"]
        self._lhs = (self.A if self.size == "sharded"
                     else self.A_unsharded.to(self.communicator.device))

    def run(self):
        import torch
        return torch.matmul(self._lhs, self.B)

    def reference_result(self):
        if self.size == "unsharded":
            return super().reference_result()
        d, r = self.communicator.world_size, self.communicator.rank
        rows = self.m // d
        return self._oracle_matmul(self.A_unsharded[r * rows:(r + 1) * rows, :], self.B)

    def comm_bytes_per_gpu(self) -> float:
        return 0.0

    def flops_per_gpu(self) -> float:
        d = self.communicator.world_size
        full = 2.0 * self.m * self.n * self.k
        return full / d if self.size == "sharded" else full
