"""Library baseline for ``tp_columnwise``: NCCL/gloo all-gather + cuBLAS matmul.

This is the path the product has to beat, kept as a side-by-side baseline
(reference ``ddlb/primitives/TPColumnwise/pytorch.py:13-104``). Differences:
``backend=gloo`` exists so the plumbing config runs on CPU, the process group
belongs to the communicator, and the reference's in-loop
``torch.cuda.empty_cache()`` is an option (``empty_cache``, default on for
fidelity) so both numbers can be reported.
"""
from __future__ import annotations

from ..options import EnvVarGuard, setup_ucc_env_vars
from .tp_columnwise import TPColumnwise


class PyTorchTPColumnwise(TPColumnwise):
    DEFAULT_OPTIONS = {"backend": "nccl", "order": "AG_before", "empty_cache": True}
    ALLOWED_VALUES = {
        "backend": ["nccl", "gloo", "ucc", "ucc/tl/nccl", "ucc/tl/cuda", "ucc/tl/ucp"],
        "order": ["AG_before", "AG_after"],
        "empty_cache": [True, False],
    }

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        import torch
        backend = self.options["backend"]
        self.env_guard = EnvVarGuard(setup_ucc_env_vars(backend))
        self.communicator.init_process_group("ucc" if backend.startswith("ucc") else backend)
        self.order = self.options["order"]
        self._flush = bool(self.options["empty_cache"]) and self.communicator.is_cuda
        out_cols = self.k if self.order == "AG_before" else self.n
        self._gathered = torch.empty(self.m, out_cols, dtype=self.A.dtype, device=self.A.device)

    def run(self):
        import torch
        import torch.distributed as dist
        if self._flush:
            torch.cuda.empty_cache()
        if self.order == "AG_before":
            dist.all_gather_into_tensor(self._gathered, self.A)
            return torch.matmul(self._gathered, self.B)
        local = torch.matmul(self.A, self.B)
        dist.all_gather_into_tensor(self._gathered, local)
        return self._gathered

    def close(self) -> None:
        self.env_guard.close()
