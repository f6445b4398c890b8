"""Compute-only ceiling for ``tp_rowwise`` (no counterpart class in the
reference, which has ``compute_only`` for columnwise only; added so the sweep
config can report the same ceiling for both primitives).

``size=sharded`` multiplies this rank's ``[m, k/d] x [k/d, n]`` partial (what a
rank really computes); ``size=unsharded`` multiplies the full ``[m, k] x [k, n]``.
"""
from __future__ import annotations

from .tp_rowwise import TPRowwise


class ComputeOnlyTPRowwise(TPRowwise):
    DEFAULT_OPTIONS = {"size": "sharded"}
    ALLOWED_VALUES = {"size": ["sharded", "unsharded"]}

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        dev = self.communicator.device
        self.size = self.options["size"]
        if self.size == "sharded":
            self._lhs, self._rhs = self.A, self.B
        else:
            self._lhs, self._rhs = self.A_unsharded.to(dev), self.B_unsharded.to(dev)

    def run(self):
        import torch
        return torch.matmul(self._lhs, self._rhs)

    def reference_result(self):
        if self.size == "unsharded":
            return self._oracle_matmul(self.A_unsharded, self.B_unsharded)
        d, r = self.communicator.world_size, self.communicator.rank
        kk = self.k // d
        return self._oracle_matmul(self.A_unsharded[:, r * kk:(r + 1) * kk],
                                   self.B_unsharded[r * kk:(r + 1) * kk, :])

    def comm_bytes_per_gpu(self) -> float:
        return 0.0

    def flops_per_gpu(self) -> float:
        d = self.communicator.world_size
        full = 2.0 * self.m * self.n * self.k
        return full / d if self.size == "sharded" else full
