#!/usr/bin/env python3
"""Round-2 A/B of the tp_rowwise schedule variants on real GPUs (launch with the in-tree
launcher, 2 or 8 ranks): production kernel vs variants 1, 2, 3, 4, 6, CUDA-event timed behind the
device barrier, max over ranks."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter  # noqa: E402


def main():
    m, n, k, dt = 8192, 1024, 8192, torch.float16
    comm = Communicator()
    comm.init_process_group()
    heap = comm.symmetric_heap(1 << 30)
    d, r = comm.world_size, comm.rank
    torch.manual_seed(42)
    a = (2 * torch.rand(m, k // d, device=comm.device) - 1).to(dt)
    b = (2 * torch.rand(k // d, n, device=comm.device) - 1).to(dt)
    for variant in (0, 1, 2, 3, 4, 6):
        op = FusedGemmReduceScatter(heap, r, d, m, n, k, dt, a, b, cta_group=2, block_n=256,
                                    variant=variant)
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
        times = comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)])
        if r == 0:
            ts = sorted(times)
            print(f"variant {variant}: median {ts[len(ts) // 2] * 1e3:.1f} us  min {ts[0] * 1e3:.1f} us", flush=True)
        op.close()
    comm.shutdown()


if __name__ == "__main__":
    main()
