"""EXPERIMENTAL symmetric heap on the CUDA virtual-memory API with an NVLS multicast window.

Drop-in for ``SymmetricHeap`` (same ``base/ptr/alloc/alloc_signals/tensor/empty/barrier``) plus
``mc_ptr(offset)``: an address whose stores reach every rank's copy (``multimem.st``) and whose
``multimem.ld_reduce`` loads return the sum over all ranks, computed in the NVSwitch. Native
part: ``csrc/vmm_heap.inl``. The allocations are exchanged between the rank processes as POSIX
file descriptors over Unix datagram sockets (``SCM_RIGHTS``).

Written after round 1's GPU budget was spent: never run on hardware. Select it with
``DDLB_SYMM_MODE=vmm`` once ``tests/test_vmm_experimental_gpu.py`` has passed.
"""
from __future__ import annotations

import ctypes
import os
import socket
import tempfile
import uuid
from typing import List, Optional

from .. import _native as N
from .symm_heap import SIGNAL_BYTES, SymmetricHeap, _ALIGN, _BARRIER_WORDS, _RawCudaBuffer


class VmmHeap(SymmetricHeap):
    def __init__(self, comm, min_bytes: int = 0):
        import torch
        self.comm = comm
        self.rank, self.world = comm.rank, comm.world_size
        self.device_index = comm.device.index
        want = max(int(min_bytes), int(os.environ.get("DDLB_SYMM_HEAP_BYTES", "0")), 64 << 20) + 2 * SIGNAL_BYTES
        self._lib = N.lib()
        self._torch = torch
        handle = ctypes.c_void_p()
        N.check(self._lib.ddlb_vmm_create(self.device_index, self.rank, self.world, want,
                                          ctypes.byref(handle)), "vmm_create")
        self._h = handle
        self.capacity = int(self._lib.ddlb_vmm_bytes(self._h))
        self._cursor = SIGNAL_BYTES
        self._sig_cursor = _BARRIER_WORDS * 4
        self._closed = False
        self._barrier_epoch = 0
        self.has_multicast = False
        if self.world > 1:
            comm.init_process_group()
            self._run_id = self._agree_run_id()
            self._map_peers()
            if self._lib.ddlb_vmm_has_multicast(self._h):
                self._setup_multicast()
            self.barrier()
            torch.cuda.synchronize()

    # ---------------------------------------------------------------- fd exchange
    def _agree_run_id(self) -> str:
        mine = uuid.uuid4().hex.encode()
        return self.comm.all_gather_bytes(mine)[0].decode()

    def _sock_path(self, tag: str, rank: int) -> str:
        base = os.path.dirname(os.environ.get("DDLB_RDZV_PREFIX", "")) or tempfile.gettempdir()
        return os.path.join(base, f"ddlb_fd_{self._run_id[:12]}_{tag}_{rank}.sock")

    def _host_barrier(self) -> None:
        import torch.distributed as dist
        dist.barrier()

    def _exchange_fds(self, fd: Optional[int], tag: str, senders: List[int]) -> List[int]:
        """Every rank in ``senders`` sends ``fd`` to all other ranks; returns fds by sender."""
        path = self._sock_path(tag, self.rank)
        rx = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
        try:
            os.unlink(path)
        except OSError:
            pass
        rx.bind(path)
        rx.settimeout(120)
        self._host_barrier()                         # every rank is bound
        out = [-1] * self.world
        try:
            if self.rank in senders:
                tx = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
                for p in range(self.world):
                    if p != self.rank:
                        socket.send_fds(tx, [bytes([self.rank])], [fd], address=self._sock_path(tag, p))
                tx.close()
            expected = len([s for s in senders if s != self.rank])
            for _ in range(expected):
                msg, fds, _flags, _addr = socket.recv_fds(rx, 16, 1)
                out[msg[0]] = fds[0]
            self._host_barrier()
        finally:
            rx.close()
            try:
                os.unlink(path)
            except OSError:
                pass
        return out

    def _map_peers(self) -> None:
        fd = ctypes.c_int(-1)
        N.check(self._lib.ddlb_vmm_export_fd(self._h, ctypes.byref(fd)), "vmm_export_fd")
        fds = self._exchange_fds(fd.value, "mem", list(range(self.world)))
        for p, peer_fd in enumerate(fds):
            if p == self.rank:
                continue
            N.check(self._lib.ddlb_vmm_import_peer(self._h, p, peer_fd), f"vmm_import_peer({p})")
            os.close(peer_fd)
        os.close(fd.value)

    def _setup_multicast(self) -> None:
        fd = ctypes.c_int(-1)
        if self.rank == 0:
            N.check(self._lib.ddlb_vmm_mc_create(self._h, ctypes.byref(fd)), "vmm_mc_create")
        fds = self._exchange_fds(fd.value if self.rank == 0 else None, "mc", [0])
        if self.rank != 0:
            N.check(self._lib.ddlb_vmm_mc_import(self._h, fds[0]), "vmm_mc_import")
            os.close(fds[0])
        else:
            os.close(fd.value)
        N.check(self._lib.ddlb_vmm_mc_add_device(self._h), "vmm_mc_add_device")
        self._host_barrier()                         # every device added before anyone binds
        N.check(self._lib.ddlb_vmm_mc_bind(self._h), "vmm_mc_bind")
        self._host_barrier()
        self.has_multicast = True

    # ------------------------------------------------------------------ addresses
    def base(self, rank: Optional[int] = None) -> int:
        ptr = self._lib.ddlb_vmm_base(self._h, self.rank if rank is None else rank)
        if not ptr:
            raise N.NativeError(f"rank {rank} is not mapped")
        return int(ptr)

    def mc_ptr(self, offset: int) -> int:
        ptr = self._lib.ddlb_vmm_mc_base(self._h)
        if not ptr:
            raise N.NativeError("no multicast window on this heap")
        return int(ptr) + offset

    def adopt_peer(self, peer_rank: int, other) -> None:
        raise NotImplementedError("single-process emulation uses SymmetricHeap")

    # -------------------------------------------------------------------- barrier
    def barrier(self, stream=None) -> None:
        if self.world == 1:
            return
        torch = self._torch
        s = stream if stream is not None else torch.cuda.current_stream()
        self._barrier_epoch += 1
        pads = (ctypes.c_void_p * self.world)(*[self.base(p) for p in range(self.world)])
        N.check(self._lib.ddlb_barrier_raw(pads, self.rank, self.world, self._barrier_epoch,
                                           ctypes.c_void_p(s.cuda_stream)), "barrier_raw")

    # ------------------------------------------------------------------ self-test
    def multicast_selftest(self, n_per_rank: int = 1024) -> bool:
        """``multimem.st`` then ``multimem.ld_reduce`` through the multicast window."""
        torch = self._torch
        if not self.has_multicast:
            raise N.NativeError("no multicast window on this heap")
        s = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        n = n_per_rank * self.world
        off = self.alloc(n * 4)                              # a data region, not the signal pads
        N.check(self._lib.ddlb_vmm_selftest_store(self._h, off, n_per_rank, s), "selftest_store")
        torch.cuda.synchronize()
        self.barrier()
        torch.cuda.synchronize()
        local = self.tensor(off, (n,), torch.float32)        # the local copy of the region
        expect = torch.arange(1, self.world + 1, device=local.device,
                              dtype=torch.float32).repeat_interleave(n_per_rank)
        stored_ok = bool(torch.equal(local, expect))
        out = torch.empty(n, dtype=torch.float32, device=local.device)
        N.check(self._lib.ddlb_vmm_selftest_reduce(self._h, off, ctypes.c_void_p(out.data_ptr()), n, s),
                "selftest_reduce")
        torch.cuda.synchronize()
        reduced_ok = bool(torch.equal(out, expect * self.world))
        self.barrier()
        torch.cuda.synchronize()
        return stored_ok and reduced_ok

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self._lib.ddlb_vmm_destroy(self._h)
