"""Symmetric heap: one device arena per rank, mapped into every peer.

Native part: ``csrc/runtime.inl`` (``cudaMalloc`` arena exported with ``cudaIpc``,
peer mapping, zeroed signal pads, device barrier kernel). The handles travel through the
communicator's key-value store (no NCCL communicator). This replaces the NCCL
communicator set-up of the reference's implementations
(``ddlb/primitives/TPColumnwise/pytorch.py:53-59``): the fused kernels address
peer memory directly as ``base[peer] + offset``.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Tuple

from .. import _native as N

SIGNAL_BYTES = 1 << 20          # must match kSignalBytes in csrc/runtime.inl
_ALIGN = 4096
_BARRIER_WORDS = 64             # words reserved at the start of the pads for the barrier

_TYPESTR = {1: "|u1", 2: "<u2", 4: "<u4", 8: "<u8"}


class _RawCudaBuffer:
    """Minimal ``__cuda_array_interface__`` carrier so torch can wrap a raw pointer."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3,
            "strides": None,
        }


class SymmetricHeap:
    """Bump allocator over the symmetric arena; identical call sequence on every rank
    yields identical offsets, which is what makes ``base[peer] + offset`` valid."""

    has_multicast = False       # the cudaIpc arena has no NVLS window (see vmm_heap.VmmHeap)

    def __init__(self, comm=None, min_bytes: int = 0, *, device: Optional[int] = None,
                 rank: Optional[int] = None, world: Optional[int] = None, exchange: bool = True):
        import torch
        self.comm = comm
        self.rank = comm.rank if comm is not None else int(rank)
        self.world = comm.world_size if comm is not None else int(world)
        self.device_index = (comm.device.index if comm is not None else int(device))
        env_bytes = int(os.environ.get("DDLB_SYMM_HEAP_BYTES", "0"))
        want = max(int(min_bytes), env_bytes, 64 << 20) + 2 * SIGNAL_BYTES
        self._lib = N.lib()
        handle = ctypes.c_void_p()
        N.check(self._lib.ddlb_symm_create(self.device_index, self.rank, self.world, want,
                                            ctypes.byref(handle)), "symm_create")
        self._h = handle
        self.capacity = int(self._lib.ddlb_symm_bytes(self._h))
        self._cursor = SIGNAL_BYTES                 # data allocations start after the pads
        self._sig_cursor = _BARRIER_WORDS * 4       # signal words after the barrier slots
        self._closed = False
        self._torch = torch
        if exchange and self.world > 1:
            self._exchange()

    # ------------------------------------------------------------- bootstrap
    def _exchange(self) -> None:
        mine = ctypes.create_string_buffer(N.IPC_HANDLE_BYTES)
        N.check(self._lib.ddlb_symm_export(self._h, mine), "symm_export")
        handles = self.comm.all_gather_bytes(mine.raw)
        blob = ctypes.create_string_buffer(b"".join(handles), N.IPC_HANDLE_BYTES * self.world)
        N.check(self._lib.ddlb_symm_open_peers(self._h, blob), "symm_open_peers")
        self.barrier()
        self._torch.cuda.synchronize()

    def adopt_peer(self, peer_rank: int, other: "SymmetricHeap") -> None:
        """Single-process emulation: use another heap of this process as a peer."""
        N.check(self._lib.ddlb_symm_adopt_peer(self._h, peer_rank, other._h), "adopt_peer")

    # ------------------------------------------------------------ allocation
    def base(self, rank: Optional[int] = None) -> int:
        ptr = self._lib.ddlb_symm_base(self._h, self.rank if rank is None else rank)
        if not ptr:
            raise N.NativeError(f"rank {rank} is not mapped")
        return int(ptr)

    def alloc(self, nbytes: int) -> int:
        """Reserve ``nbytes``; returns the offset (same on every rank)."""
        off = (self._cursor + _ALIGN - 1) // _ALIGN * _ALIGN
        if off + nbytes > self.capacity:
            raise MemoryError(
                f"symmetric heap exhausted: need {off + nbytes} B of {self.capacity} B "
                "(set DDLB_SYMM_HEAP_BYTES)")
        self._cursor = off + nbytes
        return off

    def alloc_signals(self, words: int) -> int:
        """Reserve zero-initialised uint32 signal words inside the pads; returns byte offset."""
        off = (self._sig_cursor + 255) // 256 * 256
        if off + 4 * words > SIGNAL_BYTES:
            raise MemoryError("signal pad exhausted")
        self._sig_cursor = off + 4 * words
        return off

    def mark(self) -> Tuple[int, int]:
        """Current allocation state, for ``release``."""
        return self._cursor, self._sig_cursor

    def release(self, mark: Tuple[int, int]) -> None:
        """Rewind the bump allocator to ``mark`` (collective: every rank, same point, after all
        ops that used the released memory were closed). The released signal words are zeroed
        again and a barrier separates the clearing from the next user."""
        torch = self._torch
        cursor, sig_cursor = mark
        if cursor > self._cursor or sig_cursor > self._sig_cursor:
            raise ValueError("release() takes a mark taken earlier on this heap")
        torch.cuda.synchronize()
        if self.world > 1:
            self.barrier()                       # nobody still signals into the old words
            torch.cuda.synchronize()
        if sig_cursor < self._sig_cursor:
            n = self._sig_cursor - sig_cursor
            raw = _RawCudaBuffer(self.base() + sig_cursor, n, self)
            torch.as_tensor(raw, device=f"cuda:{self.device_index}").zero_()
        self._cursor, self._sig_cursor = cursor, sig_cursor
        torch.cuda.synchronize()
        if self.world > 1:
            self.barrier()
            torch.cuda.synchronize()

    def ptr(self, offset: int, rank: Optional[int] = None) -> int:
        return self.base(rank) + offset

    def mc_ptr(self, offset: int) -> int:
        raise N.NativeError("this heap has no multicast window (use DDLB_SYMM_MODE=vmm)")

    def tensor(self, offset: int, shape: Tuple[int, ...], dtype):
        """Torch view of LOCAL heap memory at ``offset`` (no copy, no ownership)."""
        torch = self._torch
        numel = 1
        for s in shape:
            numel *= int(s)
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        raw = _RawCudaBuffer(self.base() + offset, nbytes, self)
        flat = torch.as_tensor(raw, device=f"cuda:{self.device_index}")
        return flat.view(dtype).view(*shape)

    def empty(self, shape: Tuple[int, ...], dtype):
        torch = self._torch
        numel = 1
        for s in shape:
            numel *= int(s)
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        off = self.alloc(nbytes)
        return off, self.tensor(off, shape, dtype)

    # --------------------------------------------------------------- barrier
    def barrier(self, stream=None) -> None:
        """Enqueue the device barrier kernel on ``stream`` (default: current stream)."""
        torch = self._torch
        s = stream if stream is not None else torch.cuda.current_stream()
        N.check(self._lib.ddlb_symm_barrier(self._h, ctypes.c_void_p(s.cuda_stream)), "barrier")

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self._lib.ddlb_symm_destroy(self._h)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
