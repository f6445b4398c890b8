"""Measure the NVLink path between ranks with a plain load/store kernel.

    python -m ddlb_b200.launcher -np 2 -m ddlb_b200.parallel.p2p_bench [--mib 256]

Reports GB/s for pull (read peer memory) and push (write peer memory), with only rank 0
active and with every rank active at once, for several CTA counts. These are the wire
rates the fused kernels' communication can be compared against.
"""
from __future__ import annotations

import argparse
import ctypes
import json


def main() -> None:
    import torch
    from .. import _native as N
    from ..communicator import Communicator

    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=256)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    comm = Communicator()
    comm.init_process_group()
    nbytes = args.mib << 20
    heap = comm.symmetric_heap(2 * nbytes + (64 << 20))
    off_a = heap.alloc(nbytes)
    off_b = heap.alloc(nbytes)
    lib = N.lib()
    stream = torch.cuda.current_stream()
    sp = ctypes.c_void_p(stream.cuda_stream)
    peer = (comm.rank + 1) % comm.world_size
    results = {}

    def timed(dst, src, blocks, active):
        best = 0.0
        for _ in range(args.iters):
            comm.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            heap.barrier()
            s.record()
            if active:
                N.check(lib.ddlb_copy(ctypes.c_void_p(dst), ctypes.c_void_p(src), nbytes, blocks, sp))
            e.record()
            torch.cuda.synchronize()
            if active:
                best = max(best, nbytes / (s.elapsed_time(e) * 1e6))
        return best

    for blocks in (16, 32, 64, 148, 592):
        for name, dst, src in (("pull", heap.ptr(off_a), heap.ptr(off_b, peer)),
                               ("push", heap.ptr(off_a, peer), heap.ptr(off_b))):
            uni = timed(dst, src, blocks, comm.rank == 0)
            bidi = timed(dst, src, blocks, True)
            results[f"{name}_blocks{blocks}"] = {"unidirectional_gbs": round(uni, 1),
                                                 "all_ranks_gbs": round(bidi, 1)}
    local = timed(heap.ptr(off_a), heap.ptr(off_b), 592, True)
    results["local_copy_gbs"] = round(local, 1)
    if comm.rank == 0:
        print(json.dumps({"world": comm.world_size, "mib": args.mib, "results": results}, indent=1))
    comm.shutdown()


if __name__ == "__main__":
    main()
