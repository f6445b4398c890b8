#!/usr/bin/env python3
"""How fast is the AllGather inside the fused kernel when the GEMM is negligible (n = 64)
compared with the headline n = 1024, for every transport and order? Separates "copy path" from
"contention with the GEMM's L2 traffic". Launch with the in-tree launcher:
    python -m ddlb_b200.launcher -np 2 scripts/dev/r2_ag_comm_only.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm  # noqa: E402


def timed(comm, heap, op, iters=20):
    torch.cuda.synchronize()
    heap.barrier()
    for _ in range(5):
        op.run()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        heap.barrier()
        starts[i].record()
        op.run()
        ends[i].record()
    torch.cuda.synchronize()
    times = sorted(comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)]))
    return times[len(times) // 2], times[0]


def main():
    m, k, dt = 65536, 1024, torch.bfloat16
    comm = Communicator()
    heap = comm.symmetric_heap(3 << 30)
    d, r = comm.world_size, comm.rank
    mc = getattr(heap, "has_multicast", False)
    if r == 0:
        print(f"world {d} heap {type(heap).__name__} multicast {mc}", flush=True)
    for n in ((1024,) if os.environ.get("DDLB_AB_QUICK", "0") == "1" else (64, 1024)):
        b = torch.randn(k, n, device=comm.device, dtype=dt)
        cases = [("AG_before", "pull", w) for w in (2, 4, 6)]
        cases += [("AG_before", t, w) for t in ["push"] + (["multimem"] if mc else []) for w in (2, 6)]
        cases += [("AG_after", t, 0) for t in ["tma", "st128"] + (["multimem"] if mc else [])]
        for order, transport, warps in cases:
            moved = (d - 1) / d * m * (k if order == "AG_before" else n) * 2
            op = FusedAllGatherGemm(heap, r, d, m, n, k, dt, b, order=order, transport=transport,
                                    comm_warps=warps)
            op.a_shard.normal_()
            med, best = timed(comm, heap, op)
            if r == 0:
                print(f"n={n:5d} {order:9s} transport={transport:8s} warps={warps} median {med * 1e3:7.1f} us (best "
                      f"{best * 1e3:7.1f})  inbound {moved / (med * 1e6):6.1f} GB/s  info={op.plan.info}",
                      flush=True)
            op.close()
    comm.shutdown()


if __name__ == "__main__":
    main()
