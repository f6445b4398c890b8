"""A/B of wide tiles (256 x 512 per CTA pair, two accumulators share every A stage) against the
256 x 256 tiles and cuBLAS (1 GPU). The three candidates are timed in turn inside every repetition
so that clock / power drift hits them alike; K back-to-back launches under one CUDA event pair."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", ".."))

from ddlb_b200.ops.tp_gemm import plain_gemm_plan


def once(fn, iters):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    shapes = [(65536, 1024, 1024), (8192, 8192, 2048), (8192, 2048, 4096), (4096, 4096, 4096),
              (16384, 4096, 4096), (8192, 1024, 8192), (8192, 8192, 8192), (16384, 3584, 8192),
              (16384, 8192, 3584), (16384, 16384, 16384)]
    w = torch.randn(8192, 8192, device="cuda").bfloat16()
    for _ in range(100):
        torch.matmul(w, w)
    torch.cuda.synchronize()
    del w
    for m, n, k in shapes:
        a = torch.randn(m, k, device="cuda").bfloat16()
        b = torch.randn(k, n, device="cuda").bfloat16()
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        narrow, out_n = plain_gemm_plan(a, b, wide_n=False)
        wide, out_w = plain_gemm_plan(a, b, wide_n=True)
        cands = {"narrow": narrow.run, "wide": wide.run, "cublas": lambda: torch.matmul(a, b, out=c)}
        flops = 2.0 * m * n * k
        iters = max(5, min(200, int(20e-3 / (flops / 1.4e15))))
        for fn in cands.values():
            for _ in range(5):
                fn()
        times = {name: [] for name in cands}
        for _ in range(5):
            for name, fn in cands.items():
                times[name].append(once(fn, iters))
        torch.cuda.synchronize()
        bad = int((~((out_w.float() - out_n.float()).abs() <= 1e-2 * out_n.float().abs() + 1e-2 * k ** 0.5)).sum())
        med = {name: sorted(v)[len(v) // 2] for name, v in times.items()}
        print(f"{m}x{n}x{k} | " + " | ".join(
            f"{name} {med[name]:9.2f} us {flops / med[name] / 1e6:7.0f} TF" for name in cands) +
            f" | wide/narrow {med['narrow'] / med['wide']:.3f}x wide/cublas {med['cublas'] / med['wide']:.3f}x"
            f" | mismatches {bad}", flush=True)
        narrow.close()
        wide.close()


if __name__ == "__main__":
    main()
