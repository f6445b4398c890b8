#!/usr/bin/env python3
"""Developer check: emulated multi-rank fused modes on one GPU."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200 import _native as N  # noqa: E402
from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated  # noqa: E402
from ddlb_b200.ops.tp_gemm import (FusedAllGatherGemm, FusedGemmReduceScatter,  # noqa: E402
                                   plain_gemm_plan)


def check(out, ref, k, what):
    err = (out.float() - ref).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
    bad = int((~(err <= tol)).sum())
    print(f"  {what}: max_err={err.max().item():.4g} bad={bad}/{err.numel()} nan={int(torch.isnan(out).sum())}",
          flush=True)
    return bad == 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="plain")
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--cta-group", type=int, default=1)
    ap.add_argument("--chunk-rows", type=int, default=0)
    ap.add_argument("--epochs", type=int, default=2)
    args = ap.parse_args()
    torch.manual_seed(3)
    w, g = args.world, args.cta_group
    ok = True
    try:
        if args.mode == "plain":
            a = torch.randn(1024, 256, device="cuda", dtype=torch.bfloat16)
            b = torch.randn(256, 512, device="cuda", dtype=torch.bfloat16)
            plan, out = plain_gemm_plan(a, b, cta_group=g)
            out.fill_(float("nan"))
            run_emulated([plan], 1)
            torch.cuda.synchronize()
            ok &= check(out, a.float() @ b.float(), 256, "emulated plain")
        elif args.mode in ("ag_before", "ag_after"):
            m, n, k = 512 * w, 512, 256
            rows = m // w
            a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
            b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
            ref = a.float() @ b.float()
            heaps = make_emulated_heaps(w, 2 * m * max(k, 2 * n) * 2 + (8 << 20))
            order = "AG_before" if args.mode == "ag_before" else "AG_after"
            ops = [FusedAllGatherGemm(heaps[r], r, w, m, n, k, a.dtype, b, order=order,
                                      chunk_rows=args.chunk_rows, n_comm_ctas=2 * g, cta_group=g)
                   for r in range(w)]
            print("  info", ops[0].plan.info, flush=True)
            for r, op in enumerate(ops):
                op.a_shard.copy_(a[r * rows:(r + 1) * rows])
            torch.cuda.synchronize()
            for epoch in range(1, args.epochs + 1):
                idx = epoch % len(ops[0].plans)
                for op in ops:
                    op.out[idx].fill_(float("nan"))
                run_emulated([op.plans[idx] for op in ops], epoch)
                torch.cuda.synchronize()
                for r, op in enumerate(ops):
                    ok &= check(op.out[idx], ref, k, f"{order} rank{r} epoch{epoch}")
        elif args.mode == "rs":
            m, n, k = 512 * w, 512, 128 * w
            rows, kk = m // w, k // w
            a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
            b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
            ref = a.float() @ b.float()
            heaps = make_emulated_heaps(w, 4 * m * n * 2 + (8 << 20))
            ops = [FusedGemmReduceScatter(heaps[r], r, w, m, n, k, a.dtype,
                                          a[:, r * kk:(r + 1) * kk].contiguous(),
                                          b[r * kk:(r + 1) * kk].contiguous(),
                                          chunk_rows=args.chunk_rows, cta_group=g)
                   for r in range(w)]
            print("  info", ops[0].plan.info, flush=True)
            for epoch in range(1, args.epochs + 1):
                for op in ops:
                    op.out.fill_(float("nan"))
                run_emulated([op.plan for op in ops], epoch)
                torch.cuda.synchronize()
                for r, op in enumerate(ops):
                    ok &= check(op.out, ref[r * rows:(r + 1) * rows], k, f"RS rank{r} epoch{epoch}")
    except Exception as exc:
        print(f"  EXCEPTION: {str(exc).splitlines()[0]}", flush=True)
        try:
            print("  error word:", N.read_error_word(), flush=True)
        except Exception as e2:
            print("  (error word unreadable:", str(e2).splitlines()[0], ")", flush=True)
        ok = False
    print("RESULT", args.mode, "world", w, "g", g, "OK" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
