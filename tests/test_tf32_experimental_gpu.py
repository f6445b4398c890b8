"""float32 through the kind::tf32 kernels. Written after round 1's GPU budget was spent, so
these tests have never run on hardware: they are opt-in (``DDLB_RUN_EXPERIMENTAL=1``) and the
feature itself is opt-in (``DDLB_ENABLE_TF32=1``) until they have passed on a B200."""
import os

import pytest

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(os.environ.get("DDLB_RUN_EXPERIMENTAL", "0") != "1",
                                 reason="experimental: set DDLB_RUN_EXPERIMENTAL=1")]


@pytest.fixture(autouse=True)
def _enable(monkeypatch):
    monkeypatch.setenv("DDLB_ENABLE_TF32", "1")


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("block_n", [128, 256])
@pytest.mark.parametrize("b_is_nk", [False, True])
@pytest.mark.parametrize("shape", [(256, 256, 32), (384, 320, 200), (1024, 1152, 512)])
def test_plain_gemm_tf32(cta_group, block_n, b_is_nk, shape):
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    torch.manual_seed(2)
    m, n, k = shape
    a = 2 * torch.rand(m, k, device="cuda") - 1
    b_kn = 2 * torch.rand(k, n, device="cuda") - 1
    b = b_kn.t().contiguous() if b_is_nk else b_kn
    plan, out = plain_gemm_plan(a, b, b_is_nk=b_is_nk, cta_group=cta_group, block_n=block_n)
    out.fill_(float("nan"))
    plan.run()
    torch.cuda.synchronize()
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        ref = a @ b_kn
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev
    # the reference's tolerance for float32: atol = 1e-4 * k (tp_columnwise.py:150-162)
    torch.testing.assert_close(out, ref, rtol=0, atol=1e-4 * k)


@pytest.mark.parametrize("world", [2, 4])
def test_fused_tf32_emulated(world):
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter
    torch.manual_seed(3)
    m, n, k = 512 * world, 512, 128 * world
    rows, kk = m // world, k // world
    a = 2 * torch.rand(m, k, device="cuda") - 1
    b = 2 * torch.rand(k, n, device="cuda") - 1
    ref = a.double() @ b.double()
    heaps = make_emulated_heaps(world, 8 * m * max(k, 2 * n) * 4 + (8 << 20))
    ag = [FusedAllGatherGemm(heaps[r], r, world, m, n, k, a.dtype, b) for r in range(world)]
    for r, op in enumerate(ag):
        op.a_shard.copy_(a[r * rows:(r + 1) * rows])
    rs = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype,
                                 a[:, r * kk:(r + 1) * kk].contiguous(),
                                 b[r * kk:(r + 1) * kk].contiguous()) for r in range(world)]
    torch.cuda.synchronize()
    for epoch in (1, 2):
        run_emulated([op.plan for op in ag], epoch)
        run_emulated([op.plan for op in rs], epoch)
        torch.cuda.synchronize()
        for r in range(world):
            torch.testing.assert_close(ag[r].out[0].double(), ref, rtol=0, atol=1e-4 * k)
            torch.testing.assert_close(rs[r].out.double(), ref[r * rows:(r + 1) * rows], rtol=0,
                                       atol=1e-4 * k)
