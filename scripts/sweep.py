#!/usr/bin/env python3
"""Shape sweep of both primitives in ONE process per rank (BASELINE.json config #4 and #5).

    python -m ddlb_b200.launcher -np 8 scripts/sweep.py [--sizes 1024,8192,16384] [--llama]
                                                        [--out results/sweep.csv]

For every (m, n, k): the fused kernels (tp_columnwise in both orders, tp_rowwise) next to the
library pairs they replace, NCCL + cuBLAS exactly as the reference's ``pytorch`` implementation
calls them but WITHOUT its in-loop ``empty_cache()`` (``ddlb/primitives/TPColumnwise/pytorch.py:92-102``,
``TPRowwise/pytorch.py:79-83``), and cuBLAS alone (``compute_only``). Device-generated inputs,
every fused result checked against an fp32 product, CUDA events behind the device barrier,
median over iterations of the max over ranks. The harness (``run_benchmark``) measures the same
things one subprocess per implementation; this script exists because 27 shapes x 7
implementations in subprocesses cost a quarter of an hour of an 8-GPU box.
"""
import argparse
import csv
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter  # noqa: E402
from ddlb_b200.utils.roofline import NVLINK_MEASURED_GBS, roofline_ms  # noqa: E402


def timed(comm, heap, fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(iters)]
    for i in range(iters):
        if heap is not None:
            heap.barrier()
        starts[i].record()
        fn()
        ends[i].record()
    torch.cuda.synchronize()
    ts = sorted(comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)]))
    return ts[len(ts) // 2]


def close_enough(out, ref, k):
    err = (out.float() - ref).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
    return int((~(err <= tol)).sum()) == 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1024,8192,16384")
    ap.add_argument("--llama", action="store_true", help="add the Llama-3-70B TP MLP shapes")
    ap.add_argument("--only-llama", action="store_true")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    comm = Communicator()
    d, r, dev = comm.world_size, comm.rank, comm.device
    dt = torch.bfloat16
    heap = comm.symmetric_heap(int(os.environ.get("DDLB_SYMM_HEAP_BYTES", str(7 << 30)))) if d > 1 else None
    if d > 1:
        comm.init_process_group("nccl")          # the library baselines only
    sizes = [int(x) for x in args.sizes.split(",")]
    shapes = [] if args.only_llama else [(m, n, k) for m in sizes for n in sizes for k in sizes]
    col_shapes, row_shapes = list(shapes), list(shapes)
    if args.llama or args.only_llama:
        col_shapes.append((16384, 28672, 8192))  # up projection, TP = 8 in BASELINE.json
        row_shapes.append((16384, 8192, 28672))  # down projection
    rows_out = []

    def record(prim, m, n, k, name, ms, valid, flops_gpu, comm_bytes):
        roof = roofline_ms(flops_gpu, comm_bytes, "bfloat16", link_gbs=NVLINK_MEASURED_GBS)
        rows_out.append({"primitive": prim, "m": m, "n": n, "k": k, "world": d, "implementation": name,
                         "ms": round(ms, 4), "tflops_2mnk": round(2.0 * m * n * k / (ms * 1e9), 1),
                         "pct_roofline_measured_link": round(100 * roof / ms, 1), "valid": valid})
        if r == 0:
            print(f"{prim} {m}x{n}x{k} d={d} {name:28s} {ms * 1e3:9.1f} us  "
                  f"{2.0 * m * n * k / (ms * 1e9):8.0f} TF  valid={valid}", flush=True)

    for m, n, k in col_shapes:
        if m % (256 * d):
            continue
        torch.manual_seed(11)
        a_full = (2 * torch.rand(m, k, device=dev) - 1).to(dt)
        b = (2 * torch.rand(k, n, device=dev) - 1).to(dt)
        rows = m // d
        a_sh = a_full[r * rows:(r + 1) * rows].contiguous()
        ref = a_full.float() @ b.float()
        comm_bytes = (d - 1) / d * m * k * 2
        mark = heap.mark() if heap is not None else None
        for order in ("AG_before", "AG_after"):
            if d == 1 and order == "AG_after":
                continue
            op = FusedAllGatherGemm(heap, r, d, m, n, k, dt, b, order=order)
            op.a_shard.copy_(a_sh)
            torch.cuda.synchronize()
            if heap is not None:
                heap.barrier()
            ms = timed(comm, heap, op.run, args.iters)
            ok = close_enough(op.run(), ref, k)
            torch.cuda.synchronize()
            record("tp_columnwise", m, n, k, f"fused {order}", ms, ok,
                   2.0 * m * n * k / (1 if order == "AG_before" else d), comm_bytes)
            op.close()
        if d > 1:
            gathered = torch.empty(m, k, dtype=dt, device=dev)
            out_g = torch.empty(m, n, dtype=dt, device=dev)

            def nccl_before():
                dist.all_gather_into_tensor(gathered, a_sh)
                return torch.matmul(gathered, b)

            def nccl_after():
                dist.all_gather_into_tensor(out_g, torch.matmul(a_sh, b))
                return out_g
            record("tp_columnwise", m, n, k, "nccl+cublas AG_before", timed(comm, heap, nccl_before, args.iters),
                   True, 2.0 * m * n * k, comm_bytes)
            record("tp_columnwise", m, n, k, "nccl+cublas AG_after", timed(comm, heap, nccl_after, args.iters),
                   True, 2.0 * m * n * k / d, comm_bytes)
            del gathered, out_g
        record("tp_columnwise", m, n, k, "cublas unsharded", timed(comm, heap, lambda: torch.matmul(a_full, b), args.iters),
               True, 2.0 * m * n * k, 0)
        record("tp_columnwise", m, n, k, "cublas sharded", timed(comm, heap, lambda: torch.matmul(a_sh, b), args.iters),
               True, 2.0 * m * n * k / d, 0)
        if heap is not None:
            heap.release(mark)
        del a_full, b, ref, a_sh

    for m, n, k in row_shapes:
        if m % (256 * d) or k % d:
            continue
        torch.manual_seed(12)
        kk, rows = k // d, m // d
        a_loc = (2 * torch.rand(m, kk, device=dev) - 1).to(dt)
        b_loc = (2 * torch.rand(kk, n, device=dev) - 1).to(dt)
        part = a_loc.float() @ b_loc.float()
        if d > 1:
            dist.all_reduce(part)
        ref = part[r * rows:(r + 1) * rows]
        comm_bytes = (d - 1) / d * m * n * 2
        mark = heap.mark() if heap is not None else None
        op = FusedGemmReduceScatter(heap, r, d, m, n, k, dt, a_loc, b_loc)
        torch.cuda.synchronize()
        if heap is not None:
            heap.barrier()
        ms = timed(comm, heap, op.run, args.iters)
        ok = close_enough(op.run(), ref, k)
        torch.cuda.synchronize()
        record("tp_rowwise", m, n, k, f"fused ({op.transport})", ms, ok, 2.0 * m * n * k / d, comm_bytes)
        op.close()
        if d > 1:
            shard = torch.empty(rows, n, dtype=dt, device=dev)

            def nccl_rs():
                dist.reduce_scatter_tensor(shard, torch.matmul(a_loc, b_loc))
                return shard
            record("tp_rowwise", m, n, k, "cublas+nccl reduce_scatter", timed(comm, heap, nccl_rs, args.iters),
                   True, 2.0 * m * n * k / d, comm_bytes)
            del shard
        record("tp_rowwise", m, n, k, "cublas sharded (no reduce)", timed(comm, heap, lambda: torch.matmul(a_loc, b_loc), args.iters),
               True, 2.0 * m * n * k / d, 0)
        if heap is not None:
            heap.release(mark)
        del a_loc, b_loc, part, ref

    if r == 0 and args.out and rows_out:
        os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
        with open(args.out, "w", newline="") as fh:
            w = csv.DictWriter(fh, fieldnames=list(rows_out[0].keys()))
            w.writeheader()
            w.writerows(rows_out)
    comm.shutdown()


if __name__ == "__main__":
    main()
