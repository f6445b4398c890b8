#!/usr/bin/env python3
"""Developer check of the tcgen05 GEMM variants on one GPU (numerics + timing).

    python scripts/dev_gemm_check.py --cta-group 2 --block-n 256 --b-layout kn
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

import torch  # noqa: E402

from ddlb_b200.ops.tp_gemm import plain_gemm_plan  # noqa: E402


def bench(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cta-group", type=int, default=1)
    ap.add_argument("--block-n", type=int, default=256)
    ap.add_argument("--b-layout", default="kn", choices=["kn", "nk"])
    ap.add_argument("--shapes", default="256x256x64,256x512x256,384x320x200,4096x1024x1024,65536x1024x1024")
    ap.add_argument("--dtypes", default="bfloat16,float16")
    ap.add_argument("--perf", action="store_true")
    ap.add_argument("--fp32-mode", default="3xtf32", choices=["3xtf32", "tf32"])
    ap.add_argument("--cluster-pairs", type=int, default=0)
    args = ap.parse_args()
    torch.manual_seed(0)
    ok_all = True
    for dt_name in args.dtypes.split(","):
        dt = getattr(torch, dt_name)
        for shp in args.shapes.split(","):
            m, n, k = (int(x) for x in shp.split("x"))
            a = torch.randn(m, k, device="cuda", dtype=dt)
            b_kn = torch.randn(k, n, device="cuda", dtype=dt)
            b = b_kn.t().contiguous() if args.b_layout == "nk" else b_kn
            try:
                plan, out = plain_gemm_plan(a, b, b_is_nk=args.b_layout == "nk",
                                            cta_group=args.cta_group, block_n=args.block_n,
                                            fp32_mode=args.fp32_mode, cluster_pairs=args.cluster_pairs)
            except Exception as exc:
                print(f"[{dt_name} {shp}] plan failed: {exc}")
                ok_all = False
                continue
            out.fill_(float("nan"))
            plan.run()
            torch.cuda.synchronize()
            torch.backends.cuda.matmul.allow_tf32 = False
            ref = (a.double() @ b_kn.double()).float() if dt == torch.float32 else a.float() @ b_kn.float()
            err = (out.float() - ref).abs()
            tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
            bad = int((~(err <= tol)).sum())
            status = "OK " if bad == 0 else "BAD"
            ok_all &= bad == 0
            line = (f"[{status}] g={args.cta_group} bn={args.block_n} pairs={args.cluster_pairs} b={args.b_layout} {dt_name} {shp} "
                    f"max_err={err.max().item():.4g} bad={bad}/{err.numel()} info={plan.info}")
            if bad:
                idx = (~(err <= tol)).nonzero()[:6].tolist()
                line += f" first_bad={idx} nan={int(torch.isnan(out).sum())}"
            if args.perf and bad == 0 and m * n * k >= 2 ** 30:
                t_med, t_min = bench(plan.run)
                t_ref, t_ref_min = bench(lambda: torch.matmul(a, b_kn))
                fl = 2.0 * m * n * k / 1e9
                line += (f" | ours {t_med*1e3:.1f}us ({fl/t_med:.0f} TF, best {fl/t_min:.0f})"
                         f" cublas {t_ref*1e3:.1f}us ({fl/t_ref:.0f} TF, best {fl/t_ref_min:.0f})")
                if dt == torch.float32:
                    torch.backends.cuda.matmul.allow_tf32 = True
                    t_tf, _ = bench(lambda: torch.matmul(a, b_kn))
                    torch.backends.cuda.matmul.allow_tf32 = False
                    line += f" cublas(allow_tf32) {t_tf*1e3:.1f}us ({fl/t_tf:.0f} TF)"
            print(line, flush=True)
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
