"""End-to-end stepping of a primitive with the host->device copy of the next step's
input overlapped with the current step's kernel.

Each step: copy this rank's input shard from pinned host memory into the primitive's
device input (copy stream), run the primitive (compute stream), read one result row
back to pinned host memory. ``depth`` instances of the primitive rotate so the copy
for step ``i`` can run while step ``i-1`` computes.

Why ``depth >= 3`` for multi-rank AllGather->GEMM: the input shard of an instance is
read by the PEERS' kernels (they gather it over NVLink), and a peer may lag. The
kernel of step ``j+1`` on this rank cannot finish before every peer has started its
own step ``j+1`` kernel, i.e. finished step ``j`` (the fused kernels wait for the
peers' "ready" flags). So the input used by step ``j`` may be overwritten once THIS
rank's step ``j+1`` kernel has completed; the copy for step ``i`` (same slot as step
``i - depth``) therefore waits for the completion of step ``i - depth + 1``. With
``depth = 3`` that is step ``i-2``, which leaves step ``i-1`` free to overlap.
"""
from __future__ import annotations

from typing import Callable, List, Optional


class PipelinedRunner:
    def __init__(self, impl_factory: Callable[[], object], depth: int = 3):
        import torch
        if depth < 1:
            raise ValueError("depth must be >= 1")
        self.depth = depth
        self.impls: List[object] = [impl_factory() for _ in range(depth)]
        self.comm = self.impls[0].communicator
        if self.comm.world_size > 1 and depth < 3:
            raise ValueError("multi-rank pipelining needs depth >= 3 (see module docstring)")
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
        # the slot's previous user was step i-depth; safe once step i-depth+1 has finished
        guard = i - self.depth + 1
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
