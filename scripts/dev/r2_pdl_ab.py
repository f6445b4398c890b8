"""A/B of programmatic dependent launch on the plain tcgen05 GEMM (1 GPU): K back-to-back launches
under one CUDA event pair, PDL on / off, and cuBLAS (torch.matmul into a preallocated output)."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from ddlb_b200.ops.tp_gemm import plain_gemm_plan


def timed(fn, iters=200, warmup=20, reps=5):
    for _ in range(warmup):
        fn()
    best, vals = None, []
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        vals.append(e0.elapsed_time(e1) * 1e3 / iters)
    vals.sort()
    return vals[len(vals) // 2], vals[0]


def main():
    shapes = [(65536, 1024, 1024), (4096, 1024, 1024), (8192, 1024, 8192), (1024, 1024, 1024),
              (2048, 1024, 1024), (1024, 2048, 4096), (512, 512, 8192), (16384, 8192, 1024),
              (8192, 8192, 8192)]
    w = torch.randn(8192, 8192, device="cuda").bfloat16()        # bring the clocks up first
    for _ in range(300):
        torch.matmul(w, w)
    torch.cuda.synchronize()
    for m, n, k in shapes:
        a = torch.randn(m, k, device="cuda").bfloat16()
        b = torch.randn(k, n, device="cuda").bfloat16()
        row = [f"{m}x{n}x{k}"]
        for pdl in (True, False):
            plan, out = plain_gemm_plan(a, b, pdl=pdl)
            med, best = timed(plan.run, iters=200 if m * n * k < 2 ** 38 else 30)
            row.append(f"pdl={'on ' if pdl else 'off'} {med:8.2f} us (best {best:8.2f}) "
                       f"[g={plan.info['cta_group']} bn={plan.info['block_n']}]")
            plan.close()
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        med, best = timed(lambda: torch.matmul(a, b, out=c), iters=200 if m * n * k < 2 ** 38 else 30)
        row.append(f"cublas {med:8.2f} us (best {best:8.2f})")
        print(" | ".join(row), flush=True)


if __name__ == "__main__":
    main()
