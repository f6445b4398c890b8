#!/usr/bin/env python3
"""A/B of the tp_rowwise transports on real GPUs (push / pull / multimem.ld_reduce), headline
shape and the Llama-3-70B down projection, CUDA-event timed behind the device barrier, max over
ranks. Launch: python -m ddlb_b200.launcher -np N scripts/dev/r2_rowwise_ab.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter  # noqa: E402


def main():
    comm = Communicator()
    heap = comm.symmetric_heap(3 << 30)
    d, r = comm.world_size, comm.rank
    mc = getattr(heap, "has_multicast", False)
    if r == 0:
        print(f"world {d} heap {type(heap).__name__} multicast {mc}", flush=True)
    shapes = [(8192, 1024, 8192, torch.float16)]
    if os.environ.get("DDLB_AB_LLAMA", "1") == "1":
        shapes.append((16384, 8192, 28672, torch.bfloat16))
    for m, n, k, dt in shapes:
        if k % d or m % (256 * d):
            continue
        torch.manual_seed(42)
        a = (2 * torch.rand(m, k // d, device=comm.device) - 1).to(dt)
        b = (2 * torch.rand(k // d, n, device=comm.device) - 1).to(dt)
        for transport in ["push", "pull"] + (["multimem"] if mc else []):
            for unrot in (False, True):
                if unrot and transport != "push":
                    continue
                op = FusedGemmReduceScatter(heap, r, d, m, n, k, dt, a, b, transport=transport,
                                            unrotated=unrot)
                torch.cuda.synchronize()
                heap.barrier()
                for _ in range(5):
                    op.run()
                iters = 30
                starts = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
                ends = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
                for i in range(iters):
                    heap.barrier()
                    starts[i].record()
                    op.run()
                    ends[i].record()
                torch.cuda.synchronize()
                # back to back as well: the protocol of bench.py
                s0, e0 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                heap.barrier()
                s0.record()
                for _ in range(iters):
                    op.run()
                e0.record()
                torch.cuda.synchronize()
                times = comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)])
                b2b = comm.all_reduce_max([s0.elapsed_time(e0) / iters])[0]
                if r == 0:
                    ts = sorted(times)
                    fl = 2.0 * m * n * k / 1e9
                    print(f"RS {m}x{n}x{k} transport={transport:8s} unrotated={int(unrot)}: per-iteration median "
                          f"{ts[len(ts) // 2] * 1e3:7.1f} us (min {ts[0] * 1e3:7.1f}), back-to-back "
                          f"{b2b * 1e3:7.1f} us = {fl / b2b:7.0f} TF", flush=True)
                op.close()
        del a, b
    comm.shutdown()


if __name__ == "__main__":
    main()
