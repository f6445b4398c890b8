"""Symmetric heap on the CUDA virtual-memory API with an NVLS multicast window.

Drop-in for ``SymmetricHeap`` (same ``base/ptr/alloc/alloc_signals/tensor/empty/barrier``) plus
``mc_ptr(offset)``: an address whose stores reach every rank's copy (``multimem.st``) and whose
``multimem.ld_reduce`` loads return the sum over all ranks, computed in the NVSwitch. Native
part: ``csrc/vmm_heap.inl``. The allocations are exchanged between the rank processes as POSIX
file descriptors over Unix datagram sockets (``SCM_RIGHTS``); everything else the set-up needs
(socket names, "did every rank succeed") goes through the communicator's store, so no NCCL
communicator is involved.

This is what the reference reaches through nvFuser's ``cuda`` backend with
``multicast_protocol=multimem`` (``ddlb/primitives/TPColumnwise/fuser.py:171-206``).

Set-up is collective and all-or-nothing: if ANY rank cannot create its allocation, map a peer or
join the multicast group, every rank raises ``VmmUnavailable`` (the communicator then falls back
to the cudaIpc heap in ``DDLB_SYMM_MODE=auto``); a heap without a multicast window is still a
valid heap (``has_multicast`` False, the ``multimem`` transports refuse it).
"""
from __future__ import annotations

import array
import ctypes
import os
import socket
import tempfile
import uuid
from typing import List, Optional

from .. import _native as N
from .symm_heap import SIGNAL_BYTES, SymmetricHeap, _BARRIER_WORDS


class VmmUnavailable(RuntimeError):
    """The VMM heap cannot be set up on this system (raised on every rank alike)."""


class VmmHeap(SymmetricHeap):
    def __init__(self, comm, min_bytes: int = 0):
        import torch
        self.comm = comm
        self.rank, self.world = comm.rank, comm.world_size
        self.device_index = comm.device.index
        want = max(int(min_bytes), int(os.environ.get("DDLB_SYMM_HEAP_BYTES", "0")), 64 << 20) + 2 * SIGNAL_BYTES
        self._lib = N.lib()
        self._torch = torch
        self._h = None
        self._closed = False
        self._barrier_epoch = 0
        self.has_multicast = False
        self._cursor = SIGNAL_BYTES
        self._sig_cursor = _BARRIER_WORDS * 4

        handle = ctypes.c_void_p()
        err = ""
        if self._lib.ddlb_vmm_create(self.device_index, self.rank, self.world, want,
                                     ctypes.byref(handle)) != 0:
            err = self._lib.ddlb_last_error().decode(errors="replace")
        else:
            self._h = handle
        self._agree(err, "cuMemCreate")
        self.capacity = int(self._lib.ddlb_vmm_bytes(self._h))
        if self.world > 1:
            self._run_id = comm.all_gather_bytes(uuid.uuid4().hex.encode())[0].decode()
            self._map_peers()
            want_mc = os.environ.get("DDLB_VMM_MULTICAST", "1") != "0"
            mc_err = "" if (want_mc and self._lib.ddlb_vmm_has_multicast(self._h)) else "not supported"
            if self._all_ok(mc_err):
                # a failure from here on leaves a usable heap without the multicast window
                self.has_multicast = self._all_ok(self._try(self._setup_multicast))
            self.barrier()
            torch.cuda.synchronize()

    # ------------------------------------------------------- collective agreement
    @staticmethod
    def _try(fn) -> str:
        try:
            fn()
            return ""
        except Exception as exc:                     # noqa: BLE001 - reported to every rank
            return f"{type(exc).__name__}: {exc}"[:300] or "error"

    def _all_ok(self, err: str) -> bool:
        if self.world == 1:
            return not err
        answers = self.comm.all_gather_bytes(err.encode()[:300] or b"ok")
        return all(a == b"ok" for a in answers)

    def _agree(self, err: str, what: str) -> None:
        errs = [err] if self.world == 1 else [
            b.decode(errors="replace") for b in self.comm.all_gather_bytes(err.encode()[:300] or b"ok")]
        bad = [(r, e) for r, e in enumerate(errs) if e not in ("", "ok")]
        if bad:
            self.close()
            raise VmmUnavailable(f"{what} failed on rank {bad[0][0]}: {bad[0][1]}")

    # ---------------------------------------------------------------- fd exchange
    def _sock_path(self, tag: str, rank: int) -> str:
        base = os.path.dirname(os.environ.get("DDLB_RDZV_PREFIX", "")) or tempfile.gettempdir()
        return os.path.join(base, f"ddlb_fd_{self._run_id[:12]}_{tag}_{rank}.sock")

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
        self.comm.host_barrier()                     # every rank is bound
        out = [-1] * self.world
        try:
            if self.rank in senders:
                tx = socket.socket(socket.AF_UNIX, socket.SOCK_DGRAM)
                for p in range(self.world):
                    if p != self.rank:
                        # (socket.send_fds drops its `address` argument in CPython 3.12)
                        tx.sendmsg([bytes([self.rank])],
                                   [(socket.SOL_SOCKET, socket.SCM_RIGHTS, array.array("i", [fd]))],
                                   0, self._sock_path(tag, p))
                tx.close()
            expected = len([s for s in senders if s != self.rank])
            for _ in range(expected):
                msg, fds, _flags, _addr = socket.recv_fds(rx, 16, 1)
                out[msg[0]] = fds[0]
            self.comm.host_barrier()
        finally:
            rx.close()
            try:
                os.unlink(path)
            except OSError:
                pass
        return out

    def _map_peers(self) -> None:
        """Collective: every failure is agreed on by all ranks before anyone raises."""
        fd = ctypes.c_int(-1)
        err = ""
        if self._lib.ddlb_vmm_export_fd(self._h, ctypes.byref(fd)) != 0:
            err = self._lib.ddlb_last_error().decode(errors="replace")
        self._agree(err, "cuMemExportToShareableHandle")
        fds = self._exchange_fds(fd.value, "mem", list(range(self.world)))
        for p, peer_fd in enumerate(fds):
            if p == self.rank:
                continue
            if not err and self._lib.ddlb_vmm_import_peer(self._h, p, peer_fd) != 0:
                err = f"rank {p}: " + self._lib.ddlb_last_error().decode(errors="replace")
            os.close(peer_fd)
        os.close(fd.value)
        self._agree(err, "mapping a peer allocation")

    def _setup_multicast(self) -> None:
        fd = ctypes.c_int(-1)
        err = ""
        if self.rank == 0:
            if self._lib.ddlb_vmm_mc_create(self._h, ctypes.byref(fd)) != 0:
                err = self._lib.ddlb_last_error().decode(errors="replace")
        if not self._all_ok(err):
            raise N.NativeError(f"cuMulticastCreate: {err or 'failed on rank 0'}")
        fds = self._exchange_fds(fd.value if self.rank == 0 else None, "mc", [0])
        err = ""
        if self.rank != 0:
            if self._lib.ddlb_vmm_mc_import(self._h, fds[0]) != 0:
                err = self._lib.ddlb_last_error().decode(errors="replace")
            os.close(fds[0])
        else:
            os.close(fd.value)
        if not err and self._lib.ddlb_vmm_mc_add_device(self._h) != 0:
            err = self._lib.ddlb_last_error().decode(errors="replace")
        # every device must have been added before anyone binds (and nobody may bind if one failed)
        if not self._all_ok(err):
            raise N.NativeError(f"cuMulticastAddDevice: {err or 'failed on a peer'}")
        err = ""
        if self._lib.ddlb_vmm_mc_bind(self._h) != 0:
            err = self._lib.ddlb_last_error().decode(errors="replace")
        if not self._all_ok(err):
            raise N.NativeError(f"cuMulticastBindMem: {err or 'failed on a peer'}")

    # ------------------------------------------------------------------ addresses
    def base(self, rank: Optional[int] = None) -> int:
        ptr = self._lib.ddlb_vmm_base(self._h, self.rank if rank is None else rank)
        if not ptr:
            raise N.NativeError(f"rank {rank} is not mapped")
        return int(ptr)

    def mc_ptr(self, offset: int) -> int:
        """Multicast alias of ``offset``: ``multimem.*`` instructions only."""
        ptr = self._lib.ddlb_vmm_mc_base(self._h) if self.has_multicast else None
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
            if self._h is not None:
                self._lib.ddlb_vmm_destroy(self._h)
                self._h = None
