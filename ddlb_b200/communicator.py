"""Rank / device / topology layer.

The reference's ``Communicator`` (``ddlb/communicator.py:1-80``) only parses the
launcher environment and calls ``torch.cuda.set_device``; every implementation
then creates and destroys its own process group. Here the communicator owns the
whole life cycle:

* rank / device selection (CUDA when present, otherwise a CPU mode so the
  harness, the ``gloo`` baseline and the tests run without a GPU);
* an own bootstrap: a key-value store (file store of the in-tree launcher, torchrun's agent
  store, or a TCP store hosted by rank 0) carries the handle exchange, host barriers and the
  max-over-ranks of timings, so the fused product path never creates an NCCL communicator
  (the reference boots every implementation through NCCL, ``TPColumnwise/pytorch.py:53-59``);
* a ``torch.distributed`` process group only when a library baseline asks for one
  (``pytorch;backend=nccl|gloo``);
* the native symmetric heap (peer-mapped buffers + signal pads, see
  ``ddlb_b200/parallel/symm_heap.py``) and the device-side barrier that the
  fused kernels and the timing loop use;
* an NVLink / NVSwitch topology report (``ddlb_b200/parallel/topology.py``).
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional

from . import envs


class Communicator:
    """Process-wide singleton describing this rank and its peers."""

    _instance: Optional["Communicator"] = None

    def __new__(cls, *args, **kwargs):
        if cls._instance is None:
            cls._instance = super().__new__(cls)
            cls._instance._initialized = False
        return cls._instance

    def __init__(self, device: Optional[str] = None):
        if self._initialized:
            return
        import torch

        self.rank = envs.get_rank()
        self.local_rank = envs.get_local_rank()
        self.world_size = envs.get_world_size()
        self.local_size = envs.get_local_size()

        kind = (device or envs.get_device_kind()).lower()
        if kind == "auto":
            kind = "cuda" if torch.cuda.is_available() else "cpu"
        if kind == "cuda":
            if not torch.cuda.is_available():
                raise RuntimeError("DDLB_DEVICE=cuda requested but no CUDA device is visible")
            ndev = torch.cuda.device_count()
            if self.local_size > ndev:
                raise RuntimeError(
                    f"Local size {self.local_size} exceeds available CUDA devices {ndev}")
            torch.cuda.set_device(self.local_rank)
            self.device = torch.device(f"cuda:{self.local_rank}")
        elif kind == "cpu":
            self.device = torch.device("cpu")
        else:
            raise ValueError(f"Unknown device kind {kind!r} (expected cuda, cpu or auto)")

        self._pg_owned = False
        self._symm = None
        self._store = None
        self._store_round = 0
        self._sequence = int(os.environ.get("DDLB_CHILD_SEQUENCE", "0"))
        self._initialized = True

    # ------------------------------------------------------------------ basics
    @property
    def is_cuda(self) -> bool:
        return self.device.type == "cuda"

    @property
    def default_backend(self) -> str:
        return "nccl" if self.is_cuda else "gloo"

    def __repr__(self) -> str:
        return (f"Communicator(rank={self.rank}, world_size={self.world_size}, "
                f"local_rank={self.local_rank}, local_size={self.local_size}, "
                f"device={self.device})")

    # ---------------------------------------------------------------- bootstrap
    def store(self):
        """The run's key-value store (created on first use, one per benchmark child).

        Collision-proof choices first (a guessed TCP port can be taken, and a rank that
        fails to bind leaves its peers hanging in the rendezvous):
        * one rank: an in-process ``HashStore``;
        * in-tree launcher (``DDLB_RDZV_PREFIX``): a ``FileStore`` in a fresh directory,
          one file per child sequence number;
        * torchrun (agent store on ``MASTER_PORT``): a ``PrefixStore`` over a client of
          that store, keyed by the child sequence number;
        * otherwise (mpirun / SLURM style): a ``TCPStore`` hosted by rank 0 on
          ``DDLB_MASTER_ADDR:port`` with the port rotating per child (the reference reuses
          one port, ``ddlb/envs.py:75-77``).
        """
        if self._store is not None:
            return self._store
        import datetime
        import torch.distributed as dist
        seq = self._sequence
        timeout = datetime.timedelta(seconds=int(os.environ.get("DDLB_STORE_TIMEOUT", "300")))
        prefix = os.environ.get("DDLB_RDZV_PREFIX")
        if self.world_size == 1:
            store = dist.HashStore()
        elif prefix:
            store = dist.FileStore(f"{prefix}_{seq}", self.world_size)
        elif (os.environ.get("TORCHELASTIC_USE_AGENT_STORE", "").lower() == "true"
                and os.environ.get("MASTER_PORT") and os.environ.get("MASTER_ADDR")):
            base = dist.TCPStore(os.environ["MASTER_ADDR"], int(os.environ["MASTER_PORT"]),
                                 self.world_size, is_master=False, timeout=timeout)
            run = os.environ.get("TORCHELASTIC_RUN_ID", "run")
            store = dist.PrefixStore(f"ddlb_b200/{run}/{seq}", base)
        else:
            store = dist.TCPStore(envs.get_master_addr(), envs.rendezvous_port(seq),
                                  self.world_size, is_master=(self.rank == 0), timeout=timeout,
                                  wait_for_workers=False)
        try:
            store.set_timeout(timeout)
        except Exception:
            pass
        self._store = store
        return store

    def all_gather_bytes(self, payload: bytes) -> List[bytes]:
        """One byte string from every rank, through the store (handle exchange)."""
        if self.world_size == 1:
            return [payload]
        store = self.store()
        self._store_round += 1
        tag = f"ag/{self._store_round}"
        store.set(f"{tag}/{self.rank}", payload)
        return [bytes(store.get(f"{tag}/{r}")) for r in range(self.world_size)]

    def host_barrier(self) -> None:
        """Meet the other rank PROCESSES (no device work involved)."""
        if self.world_size > 1:
            self.all_gather_bytes(b"1")

    def all_reduce_max(self, values: List[float]) -> List[float]:
        """Element-wise max over ranks (reference ``ddlb/benchmark.py:191-204``), through
        the store: it is off the measured path and needs no communicator."""
        if self.world_size == 1:
            return list(values)
        import struct
        mine = struct.pack(f"<{len(values)}d", *values)
        rows = [struct.unpack(f"<{len(values)}d", b) for b in self.all_gather_bytes(mine)]
        return [max(col) for col in zip(*rows)] if values else []

    # ----------------------------------------------------------- process group
    def init_process_group(self, backend: Optional[str] = None) -> None:
        """Create a ``torch.distributed`` process group for a LIBRARY baseline (idempotent).
        The fused implementations never call this."""
        import torch.distributed as dist

        if dist.is_initialized():
            return
        backend = backend or self.default_backend
        if backend == "nccl" and not self.is_cuda:
            raise RuntimeError("backend=nccl needs a CUDA device; use backend=gloo on CPU")
        if backend.startswith("ucc"):
            if not getattr(dist, "is_ucc_available", lambda: False)():
                raise RuntimeError(
                    "UCC is not available in this torch build "
                    "(torch.distributed.is_ucc_available() is False)")
            backend = "ucc"
        kwargs: Dict[str, Any] = {}
        if backend == "nccl":
            kwargs["device_id"] = self.device
        dist.init_process_group(backend=backend, rank=self.rank, world_size=self.world_size,
                                store=dist.PrefixStore("pg", self.store()), **kwargs)
        self._pg_owned = True

    @property
    def pg_initialized(self) -> bool:
        import torch.distributed as dist
        return dist.is_available() and dist.is_initialized()

    def barrier(self) -> None:
        """Host-visible barrier: drain this device, then meet the other ranks."""
        import torch

        if self.is_cuda:
            torch.cuda.synchronize()
        if self.world_size > 1:
            if self._symm is not None:
                # device barrier on the signal pads + host wait: no NCCL launch
                self._symm.barrier()
                torch.cuda.synchronize()
            else:
                self.host_barrier()

    # ---------------------------------------------------------- symmetric heap
    def symmetric_heap(self, min_bytes: int = 0):
        """The rank's symmetric heap (created on first use, grown never).

        ``DDLB_SYMM_MODE``: ``ipc`` - one ``cudaMalloc`` arena exported with cudaIpc;
        ``vmm`` - cuMem allocations exchanged as POSIX file descriptors, with an NVLS
        multicast window when the devices support it (required by the ``multimem``
        transports); ``auto`` (default) - ``vmm`` when every rank can set it up, else ``ipc``.
        """
        if not self.is_cuda:
            raise RuntimeError("the symmetric heap needs CUDA devices")
        if self._symm is None:
            mode = os.environ.get("DDLB_SYMM_MODE", "auto").lower()
            if mode not in ("auto", "ipc", "vmm"):
                raise ValueError(f"DDLB_SYMM_MODE must be auto, ipc or vmm (got {mode!r})")
            heap = None
            if mode in ("auto", "vmm") and self.world_size > 1:
                from .parallel.vmm_heap import VmmHeap, VmmUnavailable
                try:
                    heap = VmmHeap(self, min_bytes)
                except VmmUnavailable as exc:
                    if mode == "vmm":
                        raise
                    if self.rank == 0:
                        print(f"[ddlb_b200] VMM/NVLS heap unavailable ({exc}); using cudaIpc",
                              flush=True)
            if heap is None:
                from .parallel.symm_heap import SymmetricHeap
                heap = SymmetricHeap(self, min_bytes)
            self._symm = heap
        elif min_bytes > self._symm.capacity:
            raise RuntimeError(
                f"symmetric heap already created with {self._symm.capacity} B; "
                f"{min_bytes} B requested (set DDLB_SYMM_HEAP_BYTES)")
        return self._symm

    def topology(self) -> Dict[str, Any]:
        from .parallel.topology import topology_report
        return topology_report(self)

    # ---------------------------------------------------------------- teardown
    def shutdown(self) -> None:
        import torch.distributed as dist

        if self._symm is not None:
            try:
                self._symm.close()
            finally:
                self._symm = None
        if self._pg_owned and dist.is_available() and dist.is_initialized():
            try:
                dist.destroy_process_group()
            except Exception:
                pass
            self._pg_owned = False
        self._store = None

    @classmethod
    def reset(cls) -> None:
        """Drop the singleton (tests)."""
        if cls._instance is not None and getattr(cls._instance, "_initialized", False):
            cls._instance.shutdown()
        cls._instance = None
