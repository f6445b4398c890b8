"""Fused AG+GEMM / GEMM+RS protocols with several ranks emulated on ONE GPU.

All ranks' kernels run as one launch (``ddlb_tp_gemm_run_emulated``): their CTAs are
co-resident, so the cross-rank flag waits make progress. The "peers" are separate
arenas on the same device, so every address computation, flag, chunk order, staging
parity and epoch path of the multi-GPU kernels is exercised.
"""
import pytest

pytestmark = pytest.mark.gpu


def _tol_check(out, ref, k, what):
    err = (out.float() - ref).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
    bad = int((~(err <= tol)).sum())
    assert bad == 0, f"{what}: {bad}/{err.numel()} out of tolerance, max err {err.max().item()}"


def _inputs(world, m, n, k, dtype, rowwise):
    import torch
    torch.manual_seed(7)
    dt = getattr(torch, dtype)
    a = torch.randn(m, k, device="cuda", dtype=dt)
    b = torch.randn(k, n, device="cuda", dtype=dt)
    return a, b, a.float() @ b.float()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("order,chunk_rows,transport,unrotated", [
    ("AG_before", 0, "pull", False), ("AG_before", 256, "pull", False), ("AG_before", -1, "pull", False),
    ("AG_before", 0, "pull", True), ("AG_before", 256, "pull", True),
    ("AG_before", 0, "push", False), ("AG_before", 256, "push", False), ("AG_before", -1, "push", True),
    ("AG_after", 0, "tma", False), ("AG_after", 0, "st128", False)])
@pytest.mark.parametrize("cta_group", [1, 2])
def test_allgather_gemm_emulated(world, order, chunk_rows, transport, unrotated, cta_group):
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm
    m, n, k, dtype = 512 * world, 512, 256, "bfloat16"
    rows = m // world
    cr = rows if chunk_rows == -1 else chunk_rows          # -1: one chunk per peer (p2p_pipeline)
    a, b, ref = _inputs(world, m, n, k, dtype, False)
    heaps = make_emulated_heaps(world, 2 * m * max(k, 2 * n) * 2 + (8 << 20))
    ops = [FusedAllGatherGemm(heaps[r], r, world, m, n, k, a.dtype, b, order=order,
                              chunk_rows=cr, cta_group=cta_group, transport=transport,
                              unrotated=unrotated)
           for r in range(world)]
    for r, op in enumerate(ops):
        op.a_shard.copy_(a[r * rows:(r + 1) * rows])
        for o in op.out:
            o.fill_(float("nan"))
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):                                 # repeated launches: monotonic flags
        idx = epoch % len(ops[0].plans)
        run_emulated([op.plans[idx] for op in ops], epoch)
        torch.cuda.synchronize()
        for r, op in enumerate(ops):
            _tol_check(op.out[idx], ref, k, f"{order} rank {r} epoch {epoch}")
            if order == "AG_before":
                assert torch.equal(op.a_gathered, a), f"gathered A mismatch on rank {r}"
            op.out[idx].fill_(float("nan"))
            if order == "AG_before" and epoch < 3:
                # wipe the gathered rows of the peers: the next launch must fetch them again
                keep = op.a_shard.clone()
                op.a_gathered.fill_(float("nan"))
                op.a_shard.copy_(keep)
        torch.cuda.synchronize()
    for op in ops:
        op.close()
    for h in heaps:
        h.close()


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("chunk_rows", [0, 256])
@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("dtype,transport,unrotated", [
    ("bfloat16", "push", False), ("float16", "push", False), ("float16", "push", True),
    ("bfloat16", "pull", False), ("float16", "pull", False)])
def test_gemm_reduce_scatter_emulated(world, chunk_rows, cta_group, dtype, transport, unrotated):
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter
    m, n, k = 512 * world, 512, 128 * world
    rows, kk = m // world, k // world
    a, b, ref = _inputs(world, m, n, k, dtype, True)
    heaps = make_emulated_heaps(world, 4 * m * n * 2 + (8 << 20))
    a_sh = [a[:, r * kk:(r + 1) * kk].contiguous() for r in range(world)]
    b_sh = [b[r * kk:(r + 1) * kk].contiguous() for r in range(world)]
    ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype, a_sh[r], b_sh[r],
                                  chunk_rows=chunk_rows, cta_group=cta_group, transport=transport,
                                  unrotated=unrotated)
           for r in range(world)]
    for epoch in (1, 2, 3, 4):                              # both staging parities, twice
        for op in ops:
            op.out.fill_(float("nan"))
            # a partial that is summed before it has landed must not pass: poison the staging
            # half this epoch writes (the other half may still be read by nobody: launches of
            # one stream are serialised)
            op.staging[epoch & 1].fill_(float("nan"))
        torch.cuda.synchronize()
        run_emulated([op.plan for op in ops], epoch)
        torch.cuda.synchronize()
        for r, op in enumerate(ops):
            _tol_check(op.out, ref[r * rows:(r + 1) * rows], k, f"RS rank {r} epoch {epoch}")
    for op in ops:
        op.close()
    for h in heaps:
        h.close()


@pytest.mark.parametrize("n", [448, 320])
def test_reduce_scatter_n_not_multiple_of_block_n(n):
    """The reducers' units overhang N when an n-block does: those lanes must stay idle."""
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter
    world, m, k = 2, 512, 256
    rows, kk = m // world, k // world
    a, b, ref = _inputs(world, m, n, k, "bfloat16", True)
    heaps = make_emulated_heaps(world, 4 * m * n * 2 + (8 << 20))
    ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype,
                                  a[:, r * kk:(r + 1) * kk].contiguous(),
                                  b[r * kk:(r + 1) * kk].contiguous()) for r in range(world)]
    for epoch in (1, 2):
        for op in ops:
            op.out.fill_(float("nan"))
        run_emulated([op.plan for op in ops], epoch)
        torch.cuda.synchronize()
        for r, op in enumerate(ops):
            _tol_check(op.out, ref[r * rows:(r + 1) * rows], k, f"RS n={n} rank {r} epoch {epoch}")
    for op in ops:
        op.close()
    for h in heaps:
        h.close()


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("variant", [dict(cluster_pairs=2), dict(wide_n=True)], ids=["two_pair", "wide"])
def test_tile_variants_fused_emulated(world, variant):
    """Every fused mode on clusters of four CTAs (A multicast across two MMA pairs) and on wide
    tiles (two n-blocks per CTA pair sharing every A stage)."""
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter
    m, n, k = 512 * world, 1024, 256 * world
    rows, kk = m // world, k // world
    a, b, ref = _inputs(world, m, n, k, "bfloat16", False)
    heaps = make_emulated_heaps(world, 8 * m * max(k, 2 * n) * 2 + (8 << 20))
    for order, transport in (("AG_before", "pull"), ("AG_before", "push"), ("AG_after", "tma")):
        ops = [FusedAllGatherGemm(heaps[r], r, world, m, n, k, a.dtype, b, order=order, transport=transport,
                                  **variant) for r in range(world)]
        for r, op in enumerate(ops):
            op.a_shard.copy_(a[r * rows:(r + 1) * rows])
        torch.cuda.synchronize()
        for epoch in (1, 2):
            idx = epoch % len(ops[0].plans)
            for op in ops:
                op.out[idx].fill_(float("nan"))
            run_emulated([op.plans[idx] for op in ops], epoch)
            torch.cuda.synchronize()
            for r, op in enumerate(ops):
                _tol_check(op.out[idx], ref, k, f"{order}/{transport} {variant} rank {r} epoch {epoch}")
        for op in ops:
            op.close()
    for transport in ("push", "pull"):
        ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype,
                                      a[:, r * kk:(r + 1) * kk].contiguous(),
                                      b[r * kk:(r + 1) * kk].contiguous(), transport=transport,
                                      **variant) for r in range(world)]
        for epoch in (1, 2, 3):
            for op in ops:
                op.out.fill_(float("nan"))
                op.staging[epoch & 1].fill_(float("nan"))
            torch.cuda.synchronize()
            run_emulated([op.plan for op in ops], epoch)
            torch.cuda.synchronize()
            for r, op in enumerate(ops):
                _tol_check(op.out, ref[r * rows:(r + 1) * rows], k, f"RS/{transport} {variant} rank {r} epoch {epoch}")
        for op in ops:
            op.close()
    for h in heaps:
        h.close()


def test_symmetric_heap_views_and_signals():
    import torch
    from ddlb_b200.ops.emulation import make_emulated_heaps
    heaps = make_emulated_heaps(2, 32 << 20)
    off0, t0 = heaps[0].empty((128, 64), torch.bfloat16)
    off1, t1 = heaps[1].empty((128, 64), torch.bfloat16)
    assert off0 == off1 and off0 % 4096 == 0
    t0.fill_(3.0)
    t1.fill_(5.0)
    torch.cuda.synchronize()
    assert heaps[0].ptr(off0, 1) == heaps[1].ptr(off1)      # "peer" address arithmetic
    assert t0.float().sum().item() == 3.0 * 128 * 64 and t1.float().sum().item() == 5.0 * 128 * 64
    s0 = heaps[0].alloc_signals(100)
    s1 = heaps[0].alloc_signals(4)
    assert s1 >= s0 + 400 and s1 % 256 == 0
    with pytest.raises(MemoryError):
        heaps[0].alloc(1 << 40)
    for h in heaps:
        h.close()
