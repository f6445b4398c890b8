"""Library baseline for ``tp_rowwise``: cuBLAS matmul + NCCL/gloo reduce-scatter
(reference ``ddlb/primitives/TPRowwise/pytorch.py:13-86``)."""
from __future__ import annotations

from ..options import EnvVarGuard, setup_ucc_env_vars
from .tp_rowwise import TPRowwise


class PyTorchTPRowwise(TPRowwise):
    DEFAULT_OPTIONS = {"backend": "nccl", "empty_cache": True}
    ALLOWED_VALUES = {
        "backend": ["nccl", "gloo", "ucc", "ucc/tl/nccl", "ucc/tl/cuda", "ucc/tl/ucp"],
        "empty_cache": [True, False],
    }

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        import torch
        backend = self.options["backend"]
        self.env_guard = EnvVarGuard(setup_ucc_env_vars(backend))
        self.communicator.init_process_group("ucc" if backend.startswith("ucc") else backend)
        self._flush = bool(self.options["empty_cache"]) and self.communicator.is_cuda
        self._gloo = backend == "gloo"
        rows = self.m // self.communicator.world_size
        self._shard = torch.empty(rows, self.n, dtype=self.A.dtype, device=self.A.device)

    def run(self):
        import torch
        import torch.distributed as dist
        if self._flush:
            torch.cuda.empty_cache()
        partial = torch.matmul(self.A, self.B)
        if self._gloo:
            # gloo has no reduce_scatter: all-reduce, then keep this rank's rows
            dist.all_reduce(partial)
            rows = self._shard.shape[0]
            r = self.communicator.rank
            self._shard.copy_(partial[r * rows:(r + 1) * rows])
        else:
            dist.reduce_scatter_tensor(self._shard, partial, op=dist.ReduceOp.SUM)
        return self._shard

    def close(self) -> None:
        self.env_guard.close()
