"""Experimental tp_rowwise schedule variants (separate kernel instantiations; the production
kernels are unchanged). Written after round 1's GPU budget was spent: opt-in until they have
run on hardware (``DDLB_RUN_EXPERIMENTAL=1``)."""
import os

import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("DDLB_RUN_EXPERIMENTAL", "0") != "1",
                                 reason="experimental: set DDLB_RUN_EXPERIMENTAL=1")]


@pytest.mark.parametrize("variant", [1, 2, 3, 4, 6])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_rowwise_variants_emulated(variant, world):
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter
    torch.manual_seed(9)
    m, n, k = 512 * world, 512, 128 * world
    rows, kk = m // world, k // world
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    ref = a.float() @ b.float()
    heaps = make_emulated_heaps(world, 4 * m * n * 2 + (8 << 20))
    ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype,
                                  a[:, r * kk:(r + 1) * kk].contiguous(),
                                  b[r * kk:(r + 1) * kk].contiguous(), cta_group=2, block_n=256,
                                  variant=variant) for r in range(world)]
    for epoch in (1, 2, 3):
        for op in ops:
            op.out.fill_(float("nan"))
        run_emulated([op.plan for op in ops], epoch)
        torch.cuda.synchronize()
        for r, op in enumerate(ops):
            err = (op.out.float() - ref[r * rows:(r + 1) * rows]).abs()
            tol = 2e-2 * (k ** 0.5) + 1e-2 * ref[r * rows:(r + 1) * rows].abs()
            assert int((~(err <= tol)).sum()) == 0, f"variant {variant} rank {r} epoch {epoch}"


@pytest.mark.parametrize("world", [2, 4])
def test_ag_after_register_push_variant_emulated(world):
    """AG_after with peers written by 128-bit stores from the staged rows (variant 4)."""
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm
    torch.manual_seed(10)
    m, n, k = 512 * world, 512, 256
    rows = m // world
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    ref = a.float() @ b.float()
    heaps = make_emulated_heaps(world, 8 * m * n * 2 + (8 << 20))
    ops = [FusedAllGatherGemm(heaps[r], r, world, m, n, k, a.dtype, b, order="AG_after",
                              cta_group=2, block_n=256, variant=4) for r in range(world)]
    for r, op in enumerate(ops):
        op.a_shard.copy_(a[r * rows:(r + 1) * rows])
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):
        idx = epoch % 2
        for op in ops:
            op.out[idx].fill_(float("nan"))
        run_emulated([op.plans[idx] for op in ops], epoch)
        torch.cuda.synchronize()
        for r, op in enumerate(ops):
            err = (op.out[idx].float() - ref).abs()
            tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
            assert int((~(err <= tol)).sum()) == 0, f"rank {r} epoch {epoch}"
