#!/usr/bin/env python3
"""Device timeline of one fused AG->GEMM launch on real GPUs (pull transport): when do the comm
warps start, how long does a 32 KiB unit take, when do the GEMM tiles get their chunks?
Launch: python -m ddlb_b200.launcher -np 2 scripts/dev/r2_trace_ag.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter  # noqa: E402
from ddlb_b200.utils.trace import decode, summarize  # noqa: E402


def main():
    m, k, dt = 65536, 1024, torch.bfloat16
    comm = Communicator()
    heap = comm.symmetric_heap(3 << 30)
    d, r = comm.world_size, comm.rank
    b2b = int(os.environ.get("DDLB_TRACE_B2B", "10"))
    for n, warps in ((1024, 2),):
        b = torch.randn(k, n, device=comm.device, dtype=dt)
        op = FusedAllGatherGemm(heap, r, d, m, n, k, dt, b, transport="pull", comm_warps=warps,
                                trace=True, block_n=256, cta_group=2)
        op.a_shard.normal_()
        torch.cuda.synchronize()
        heap.barrier()
        for _ in range(3):
            op.run()
        torch.cuda.synchronize()
        op.trace_buf.zero_()
        heap.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(b2b):                       # the buffer keeps the LAST launch's timeline
            op.run()
        e.record()
        torch.cuda.synchronize()
        ev = decode(op.trace_buf)
        if r == 0:
            print(f"=== rank {r} n={n} comm_warps={warps}: {b2b} back-to-back launches, "
                  f"{s.elapsed_time(e) * 1e3 / b2b:.1f} us each (event pair)")
            print(summarize(ev), flush=True)
        if r == 0:
            print(summarize(ev))
            # per-unit duration of the traced comm warps (warp 2 lane 0 of every CTA)
            per_cta = {}
            for x in ev:
                if x["role"] == "comm" and x["tag"] in ("flag_waited", "unit_ready", "unit_done"):
                    per_cta.setdefault(x["cta"], []).append((x["t_ns"], x["tag"]))
            durs, fences = [], []
            for seq in per_cta.values():
                seq.sort()
                prev = None
                last_ready = 0
                for t, tag in seq:
                    if tag == "unit_ready" and prev is not None:
                        durs.append(t - prev)
                    if tag == "unit_done":
                        fences.append(t - last_ready)
                        prev = t
                    if tag == "flag_waited":
                        prev = t
                    if tag == "unit_ready":
                        last_ready = t
            durs.sort(); fences.sort()
            if durs:
                print(f"unit copy (32 KiB, one warp): median {durs[len(durs)//2]/1e3:.2f} us, p10 {durs[len(durs)//10]/1e3:.2f}, "
                      f"p90 {durs[9*len(durs)//10]/1e3:.2f}; fence+announce median {fences[len(fences)//2]/1e3:.2f} us", flush=True)
        op.close()
    # ---- tp_rowwise headline, push transport: tiles vs reducer units
    m2, n2, k2, dt2 = 8192, 1024, 8192, torch.float16
    a = (2 * torch.rand(m2, k2 // d, device=comm.device) - 1).to(dt2)
    b = (2 * torch.rand(k2 // d, n2, device=comm.device) - 1).to(dt2)
    for transport in ("push",) + (("multimem",) if getattr(heap, "has_multicast", False) else ()):
        op = FusedGemmReduceScatter(heap, r, d, m2, n2, k2, dt2, a, b, transport=transport, trace=True,
                                    cta_group=2, block_n=256)
        torch.cuda.synchronize()
        heap.barrier()
        for _ in range(3):
            op.run()
        torch.cuda.synchronize()
        op.trace_buf.zero_()
        heap.barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(b2b):
            op.run()
        e.record()
        torch.cuda.synchronize()
        ev = decode(op.trace_buf)
        if r == 0:
            print(f"=== RS {transport} rank {r}: {b2b} back-to-back launches, {s.elapsed_time(e) * 1e3 / b2b:.1f} us each")
            print(summarize(ev), flush=True)
        op.close()
    comm.shutdown()


if __name__ == "__main__":
    main()
