"""End-to-end stepping of a primitive with the host->device copy of the next step's
input overlapped with the current step's kernel.

Each step: copy this rank's input shard from pinned host memory into the primitive's
device input (copy stream), run the primitive (compute stream), read one result row
back to pinned host memory. ``depth`` instances of the primitive rotate so the copy
for step ``i`` can run while step ``i-1`` computes.

Why ``depth = 2`` is enough, also for multi-rank AllGather->GEMM: the input shard of an
instance is read by the PEERS' kernels (they gather it over NVLink), but the fused kernel
returns only after every peer has finished reading it (closing handshake in
``csrc/tp_gemm_sm100.cuh``), exactly like ``ncclAllGather`` and its send buffer. The copy for
step ``i`` (same slot as step ``i - depth``) therefore waits for THIS rank's step
``i - depth`` kernel and nothing else.
"""
from __future__ import annotations

from typing import Callable, List, Optional


class PipelinedRunner:
    def __init__(self, impl_factory: Callable[[], object], depth: int = 2):
        import torch
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.depth = depth
        self.impls: List[object] = [impl_factory() for _ in range(depth)]
        self.comm = self.impls[0].communicator
        self.copy_stream = torch.cuda.Stream()
        self._done: List[Optional[object]] = []      # event per step: kernel finished
        self._step = 0
        first = self.impls[0]
        self.input_shape = tuple(first.A.shape)
        self.input_dtype = first.A.dtype

    def step(self, host_in, host_out=None):
        """One end-to-end step. ``host_in``: pinned tensor shaped like the rank's input
        shard. ``host_out``: optional pinned 1-D tensor receiving row 0 of the result."""
        import torch
        i = self._step
        impl = self.impls[i % self.depth]
        compute = torch.cuda.current_stream()
        # the slot's previous user was step i-depth: wait for that kernel
        guard = i - self.depth
        if guard >= 0 and guard < len(self._done):
            self.copy_stream.wait_event(self._done[guard])
        with torch.cuda.stream(self.copy_stream):
            impl.A.copy_(host_in, non_blocking=True)
            copied = torch.cuda.Event()
            copied.record(self.copy_stream)
        compute.wait_event(copied)
        out = impl.run()
        done = torch.cuda.Event()
        done.record(compute)
        self._done.append(done)
        if host_out is not None:
            host_out.copy_(out[0], non_blocking=True)
        self._step += 1
        return out

    @property
    def gpu_launches(self) -> int:
        return sum(getattr(impl, "gpu_launches", 0) for impl in self.impls)

    def close(self) -> None:
        import torch
        torch.cuda.synchronize()
        for impl in self.impls:
            impl.close()
        self.impls = []
        self._done = []
