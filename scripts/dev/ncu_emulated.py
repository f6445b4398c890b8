#!/usr/bin/env python3
"""One emulated multi-rank launch of a fused kernel on ONE GPU, for `ncu` (which cannot wrap a
multi-process run). Flags are epoch-valued and ncu restores memory between its replay passes,
so every pass sees the launch from the same state.  usage: ncu_emulated.py {ag|rs} [world]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch  # noqa: E402

from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter  # noqa: E402


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "ag"
    world = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    torch.manual_seed(0)
    dt = torch.bfloat16
    if mode == "ag":
        m, n, k = 8192 * world, 1024, 1024
        rows = m // world
        a = torch.randn(m, k, device="cuda", dtype=dt)
        b = torch.randn(k, n, device="cuda", dtype=dt)
        heaps = make_emulated_heaps(world, 2 * m * max(k, 2 * n) * 2 + (64 << 20))
        ops = [FusedAllGatherGemm(heaps[r], r, world, m, n, k, dt, b, order="AG_before") for r in range(world)]
        for r, op in enumerate(ops):
            op.a_shard.copy_(a[r * rows:(r + 1) * rows])
        plans = [op.plan for op in ops]
        ref, outs = a.float() @ b.float(), [op.out[0] for op in ops]
    else:
        m, n, k = 8192, 1024, 4096 * world
        rows, kk = m // world, k // world
        a = torch.randn(m, k, device="cuda", dtype=dt)
        b = torch.randn(k, n, device="cuda", dtype=dt)
        heaps = make_emulated_heaps(world, 4 * m * n * 2 + (64 << 20))
        ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, dt, a[:, r * kk:(r + 1) * kk].contiguous(),
                                      b[r * kk:(r + 1) * kk].contiguous()) for r in range(world)]
        plans = [op.plan for op in ops]
        full = a.float() @ b.float()
        ref, outs = None, [op.out for op in ops]
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):
        run_emulated(plans, epoch)
    torch.cuda.synchronize()
    for r, out in enumerate(outs):
        want = ref if mode == "ag" else full[r * rows:(r + 1) * rows]
        err = (out.float() - want).abs()
        assert int((~(err <= 2e-2 * (k ** 0.5) + 1e-2 * want.abs())).sum()) == 0, f"rank {r} wrong"
    print(f"{mode} world={world} ok")


if __name__ == "__main__":
    main()
