"""Run several ranks' fused plans as ONE launch on one GPU (tests / CI on 1 GPU).

Kernels of different ranks wait on each other's flags; launched separately on one
GPU nothing guarantees they are resident together. ``ddlb_tp_gemm_run_emulated``
splits one grid between the ranks instead (``tp_gemm_emulated_kernel``).
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence

from .. import _native as N
from ..parallel.symm_heap import SymmetricHeap


def make_emulated_heaps(world: int, nbytes: int, device: int = 0) -> List[SymmetricHeap]:
    heaps = [SymmetricHeap(None, nbytes, device=device, rank=r, world=world, exchange=False)
             for r in range(world)]
    for r, h in enumerate(heaps):
        for p, other in enumerate(heaps):
            if p != r:
                h.adopt_peer(p, other)
    return heaps


def run_emulated(plans: Sequence, epoch: int, stream=None) -> None:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    arr = (ctypes.c_void_p * len(plans))(*[p._h for p in plans])
    N.check(N.lib().ddlb_tp_gemm_run_emulated(arr, len(plans), epoch,
                                              ctypes.c_void_p(s.cuda_stream)), "run_emulated")
    for p in plans:
        p.epoch = epoch
        p.launches += 1
