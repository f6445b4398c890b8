"""tcgen05 GEMM kernel vs a plain PyTorch fp32 reference of the same op (1 GPU)."""
import pytest

pytestmark = pytest.mark.gpu


def _check(out, ref, k):
    import torch
    err = (out.float() - ref).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
    bad = int((~(err <= tol)).sum())
    assert bad == 0, f"{bad}/{err.numel()} elements out of tolerance, max err {err.max().item()}"


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("block_n", [128, 256])
@pytest.mark.parametrize("b_is_nk", [False, True])
@pytest.mark.parametrize("dtype", ["bfloat16", "float16"])
@pytest.mark.parametrize("shape", [(256, 256, 64), (384, 320, 200), (1024, 1152, 512)])
def test_plain_gemm(cta_group, block_n, b_is_nk, dtype, shape):
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    from ddlb_b200 import _native as N
    torch.manual_seed(1)
    m, n, k = shape
    dt = getattr(torch, dtype)
    a = torch.randn(m, k, device="cuda", dtype=dt)
    b_kn = torch.randn(k, n, device="cuda", dtype=dt)
    b = b_kn.t().contiguous() if b_is_nk else b_kn
    plan, out = plain_gemm_plan(a, b, b_is_nk=b_is_nk, cta_group=cta_group, block_n=block_n)
    out.fill_(float("nan"))
    plan.run()
    plan.run()                      # back-to-back launches reuse barriers/TMEM correctly
    torch.cuda.synchronize()
    assert N.loaded()
    _check(out, a.float() @ b_kn.float(), k)


def test_gemm_matches_reference_tolerance_headline_shape():
    """Uniform [-1,1) inputs and the reference's atol = 1e-3 * k."""
    import torch
    from ddlb_b200.ops.tp_gemm import gemm
    torch.manual_seed(42)
    m, n, k = 8192, 1024, 1024
    a = (2 * torch.rand(m, k, device="cuda") - 1).bfloat16()
    b = (2 * torch.rand(k, n, device="cuda") - 1).bfloat16()
    out = gemm(a, b)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.float(), a.float() @ b.float(), rtol=0, atol=1e-3 * k)


def test_unsupported_shapes_fail_loudly():
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    from ddlb_b200._native import NativeError
    a = torch.randn(128, 64, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(64, 100, device="cuda", dtype=torch.bfloat16)    # n % 64 != 0
    with pytest.raises(NativeError):
        plain_gemm_plan(a, b)
    with pytest.raises(TypeError):
        plain_gemm_plan(a.double(), b.double())
    # float32 is served (3xTF32) through the [n,k] operand layout; a raw [k,n] config is refused
    from ddlb_b200 import _native as N
    import ctypes
    cfg = N.TpGemmConfig()
    cfg.mode, cfg.dtype, cfg.b_layout, cfg.world = N.MODE_PLAIN, N.DTYPE_F32, N.B_KN, 1
    cfg.m, cfg.n, cfg.k = 128, 128, 64
    assert N.lib().ddlb_tp_gemm_sig_words(ctypes.byref(cfg)) < 0
    assert "[n,k]" in N.lib().ddlb_last_error().decode()


@pytest.mark.parametrize("b_is_nk", [False, True])
@pytest.mark.parametrize("shape", [(256, 512, 64), (1024, 1024, 1024), (2304, 1536, 4096), (8192, 2048, 512),
                                   (512, 1000, 200)])
def test_two_pair_clusters_multicast_a(shape, b_is_nk):
    """Clusters of four CTAs: two MMA pairs on neighbouring n-blocks, A multicast to both."""
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    torch.manual_seed(4)
    m, n, k = shape
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    b_kn = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    if not b_is_nk and n % 64:
        pytest.skip("B [k,n] needs n % 64 == 0")
    b = b_kn.t().contiguous() if b_is_nk else b_kn
    plan, out = plain_gemm_plan(a, b, b_is_nk=b_is_nk, cta_group=2, block_n=256, cluster_pairs=2)
    assert plan.info["cluster"] == 4 and plan.info["pairs"] == 2 and plan.info["grid"] % 4 == 0
    for _ in range(2):
        out.fill_(float("nan"))
        plan.run()
        torch.cuda.synchronize()
        ref = a.float() @ b_kn.float()
        err = (out.float() - ref).abs()
        tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
        assert int((~(err <= tol)).sum()) == 0, f"max err {err.max().item()}"


@pytest.mark.parametrize("pdl", [True, False])
def test_programmatic_dependent_launch_orders_memory(pdl):
    """Back-to-back launches with programmatic stream serialization: the kernel's prologue may run
    early, but every read must see what the previous kernel of the stream wrote, whether that
    kernel is a PyTorch one (the input changes before every launch) or our own (a chain of two
    GEMMs where the second consumes the first one's output)."""
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    torch.manual_seed(7)
    m, n, k = 4096, 1024, 1024
    a = torch.empty(m, k, device="cuda", dtype=torch.bfloat16)
    b1 = (torch.randn(k, n, device="cuda") / k ** 0.5).bfloat16()
    b2 = (torch.randn(n, n, device="cuda") / n ** 0.5).bfloat16()
    plan1, h = plain_gemm_plan(a, b1, pdl=pdl)
    plan2, out = plain_gemm_plan(h, b2, pdl=pdl)
    inputs = [torch.randn(m, k, device="cuda").bfloat16() for _ in range(6)]
    results = []
    for x in inputs:                      # nothing synchronises inside this loop
        a.copy_(x)
        plan1.run()
        plan2.run()
        results.append(out.clone())
        out.fill_(float("nan"))
        h.fill_(float("nan"))
    torch.cuda.synchronize()
    for x, got in zip(inputs, results):
        h_ref = (x.float() @ b1.float()).bfloat16()
        _check(got, h_ref.float() @ b2.float(), n)


@pytest.mark.parametrize("shape", [(1024, 1024, 1024), (2048, 1024, 512), (384, 320, 200), (4096, 1024, 256)])
def test_auto_tile_shape(shape):
    """Tile shape left on auto: small problems resolve to BLOCK_N 128 (one wave, more SMs busy)."""
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    torch.manual_seed(3)
    m, n, k = shape
    a = torch.randn(m, k, device="cuda").bfloat16()
    b = torch.randn(k, n, device="cuda").bfloat16()
    plan, out = plain_gemm_plan(a, b)
    want = 128 if ((m + 255) // 256) * ((n + 127) // 128) <= plan.info["grid"] // 2 else 256
    assert plan.info["block_n"] == want and plan.info["cta_group"] == 2
    out.fill_(float("nan"))
    plan.run()
    plan.run()
    torch.cuda.synchronize()
    _check(out, a.float() @ b.float(), k)


@pytest.mark.parametrize("b_is_nk", [False, True])
@pytest.mark.parametrize("shape", [(512, 512, 256), (384, 960, 200), (1024, 1536, 1024), (2048, 1024, 4096)])
def test_wide_tiles(shape, b_is_nk):
    """256 x 512 tiles per CTA pair: two accumulators share every A stage (explicit request only)."""
    import torch
    from ddlb_b200.ops.tp_gemm import plain_gemm_plan
    torch.manual_seed(5)
    m, n, k = shape
    a = torch.randn(m, k, device="cuda").bfloat16()
    b_kn = torch.randn(k, n, device="cuda").bfloat16()
    b = b_kn.t().contiguous() if b_is_nk else b_kn
    plan, out = plain_gemm_plan(a, b, b_is_nk=b_is_nk, cta_group=2, block_n=256,
                                wide_n=True)
    assert plan.info["wide"] == 1 and plan.info["stages"] == 4
    out.fill_(float("nan"))
    for _ in range(3):
        plan.run()
    torch.cuda.synchronize()
    _check(out, a.float() @ b_kn.float(), k)
    with pytest.raises(Exception, match="wide_n"):
        plain_gemm_plan(a[:, :64].contiguous(), b_kn[:64, :768].contiguous(), wide_n=True)   # 3 n-blocks
