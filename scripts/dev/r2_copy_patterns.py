#!/usr/bin/env python3
"""Which access pattern fills NVLink from inside a kernel? Pull (and push) 128 MiB from the next
rank with every rank active, for the patterns / launch shapes the fused kernels can use.
Launch: python -m ddlb_b200.launcher -np 2 scripts/dev/r2_copy_patterns.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200 import _native as N  # noqa: E402
from ddlb_b200.communicator import Communicator  # noqa: E402


def main():
    comm = Communicator()
    nbytes = 128 << 20
    heap = comm.symmetric_heap(2 * nbytes + (64 << 20))
    off_a, off_b = heap.alloc(nbytes), heap.alloc(nbytes)
    lib = N.lib()
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    peer = (comm.rank + 1) % comm.world_size
    if comm.rank == 0:
        print(f"world {comm.world_size} heap {type(heap).__name__}", flush=True)
    big = 226 * 1024
    # (pattern, unit, smem carve-out, flags): does the GEMM's shared-memory footprint (tiny L1) or
    # the per-unit fence explain the in-kernel rate, and do L1-bypassing loads / prefetch fix it?
    cases = [(1, 32768, 0, 0), (1, 32768, big, 0), (1, 32768, big, 1), (1, 32768, big, 2), (1, 32768, big, 3),
             (1, 32768, big, 6), (1, 32768, big, 7), (1, 32768, 0, 2), (3, 32768, big, 0)]
    shapes = [(148, 64), (148, 128), (148, 192), (148, 384)]
    dst, src = heap.ptr(off_a), heap.ptr(off_b, peer)
    for pattern, unit, smem, flags in cases:
        for blocks, threads in shapes:
            best = 0.0
            for _ in range(4):
                comm.barrier()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                heap.barrier()
                s.record()
                N.check(lib.ddlb_copy_pattern(ctypes.c_void_p(dst), ctypes.c_void_p(src), nbytes, blocks,
                                              threads, pattern, unit, smem, flags, sp), "copy_pattern")
                e.record()
                torch.cuda.synchronize()
                best = max(best, nbytes / (s.elapsed_time(e) * 1e6))
            worst = min(comm.all_reduce_max([-best]))
            if comm.rank == 0:
                print(f"pull pattern={pattern} unit={unit:6d} smem={smem // 1024:3d}K flags={flags} "
                      f"grid={blocks:3d}x{threads:3d}: {-worst:6.1f} GB/s per GPU (all ranks active)", flush=True)
    comm.shutdown()


if __name__ == "__main__":
    main()
