"""End-to-end stepping of a primitive with the host->device copy of the next step's
input overlapped with the current step's kernel.

Each step: copy this rank's input shard from pinned host memory into the primitive's
device input (copy stream), run the primitive (compute stream), read the result (or a
row slice of it) back to pinned host memory (read stream). ``depth`` instances of the
primitive rotate so the copy for step ``i`` and the read-back of step ``i-2`` can run
while step ``i-1`` computes; PCIe is full duplex, so a step costs
``max(H2D, kernel, D2H)`` instead of their sum.

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
        self.copy_stream = torch.cuda.Stream()       # host -> device
        self.read_stream = torch.cuda.Stream()       # device -> host
        self._done: List[Optional[object]] = []      # event per step: kernel finished
        self._read: List[Optional[object]] = []      # event per step: result read back
        self._step = 0
        first = self.impls[0]
        self.input_shape = tuple(first.A.shape)
        self.input_dtype = first.A.dtype

    def step(self, host_in, host_out=None, rows: Optional[slice] = None):
        """One end-to-end step. ``host_in``: pinned tensor shaped like the rank's input shard.
        ``host_out``: optional pinned tensor receiving ``result[rows]`` (default: row 0)."""
        import torch
        i = self._step
        impl = self.impls[i % self.depth]
        compute = torch.cuda.current_stream()
        prev = i - self.depth                        # the slot's previous user
        if 0 <= prev < len(self._done):
            self.copy_stream.wait_event(self._done[prev])     # its kernel has read the input
        with torch.cuda.stream(self.copy_stream):
            impl.A.copy_(host_in, non_blocking=True)
            copied = torch.cuda.Event()
            copied.record(self.copy_stream)
        compute.wait_event(copied)
        if 0 <= prev < len(self._read) and self._read[prev] is not None:
            compute.wait_event(self._read[prev])              # its result has left the device
        out = impl.run()
        done = torch.cuda.Event()
        done.record(compute)
        self._done.append(done)
        read = None
        if host_out is not None:
            self.read_stream.wait_event(done)
            with torch.cuda.stream(self.read_stream):
                host_out.copy_(out[rows] if rows is not None else out[0], non_blocking=True)
                read = torch.cuda.Event()
                read.record(self.read_stream)
        self._read.append(read)
        self._step += 1
        return out

    def drain(self) -> None:
        """Make the current stream wait for every copy issued so far (end of a timed region)."""
        import torch
        compute = torch.cuda.current_stream()
        for ev in self._read[-self.depth:]:
            if ev is not None:
                compute.wait_event(ev)

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
        self._read = []
