#!/usr/bin/env python3
"""Round-2 experiment: how fast is the AllGather inside the fused kernel when the GEMM is
negligible (n = 64) compared with the headline n = 1024? Separates "copy path" from
"contention with the GEMM's L2 traffic" (ROADMAP section 2). Launch with the in-tree launcher."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm  # noqa: E402


def main():
    m, k, dt = 65536, 1024, torch.bfloat16
    comm = Communicator()
    comm.init_process_group()
    heap = comm.symmetric_heap(3 << 30)
    d, r = comm.world_size, comm.rank
    moved = (d - 1) / d * m * k * 2
    for n in (64, 1024):
        b = torch.randn(k, n, device=comm.device, dtype=dt)
        for transport in ("cta_pull", "cta", "cta_tma"):
            op = FusedAllGatherGemm(heap, r, d, m, n, k, dt, b, transport=transport)
            op.a_shard.normal_()
            torch.cuda.synchronize()
            heap.barrier()
            for _ in range(5):
                op.run()
            iters = 20
            starts = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
            ends = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
            for i in range(iters):
                heap.barrier()
                starts[i].record()
                op.run()
                ends[i].record()
            torch.cuda.synchronize()
            times = sorted(comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)]))
            if r == 0:
                med = times[len(times) // 2]
                print(f"n={n:5d} transport={transport:8s} median {med * 1e3:7.1f} us  "
                      f"gather rate {moved / (med * 1e6):6.1f} GB/s  info={op.plan.info}", flush=True)
            op.close()
    comm.shutdown()


if __name__ == "__main__":
    main()
