"""``ddlb_tp_gemm_run_emulated_static``: the emulated multi-rank launch for hosts that may
not allocate device memory while running (the whole device already belongs to their own
arena, or their loop must never wait on an allocator). Its scratch is a fixed array inside
the library instead of a per-call ``cudaMallocAsync``.

The entry point is called straight through the C ABI (ctypes) and its results are compared
with those of the allocating entry point ``ddlb_tp_gemm_run_emulated`` on the same plans,
twice in a row. The scratch is shared by every call and keeps what the last call left in it,
so between the two calls a DIFFERENT set of plans (other shape, four ranks instead of two)
goes through the same entry point: the second call must not see any of that. The amount of
free device memory is sampled around the calls to check the claim in the name. Kept in a file
that pytest collects last so a ``-x`` run has exercised everything else first.
"""
import ctypes

import pytest

pytestmark = pytest.mark.gpu


def _within_tolerance(got, want, k):
    err = (got.float() - want.float()).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * want.float().abs()
    return int((~(err <= tol)).sum()), err.max().item()


def test_static_scratch_entry_point_matches_allocating_one():
    import torch
    from ddlb_b200 import _native as N
    from ddlb_b200.ops.emulation import make_emulated_heaps
    from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter

    lib = N.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    world, m, n, k = 2, 1024, 512, 512
    rows, kk = m // world, k // world
    torch.manual_seed(11)
    a = torch.randn(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(k, n, device="cuda", dtype=torch.bfloat16)
    exact = a.float() @ b.float()
    heaps = make_emulated_heaps(world, 32 << 20)
    ops = [FusedGemmReduceScatter(heaps[r], r, world, m, n, k, a.dtype,
                                  a[:, r * kk:(r + 1) * kk].contiguous(),
                                  b[r * kk:(r + 1) * kk].contiguous()) for r in range(world)]
    plans = (ctypes.c_void_p * world)(*[op.plan._h for op in ops])

    # existing entry point (allocates its parameter scratch): the expected output
    N.check(lib.ddlb_tp_gemm_run_emulated(plans, world, 1, stream), "run_emulated")
    torch.cuda.synchronize()
    expected = [op.out.clone() for op in ops]
    for r in range(world):
        bad, worst = _within_tolerance(expected[r], exact[r * rows:(r + 1) * rows], k)
        assert bad == 0, f"allocating entry point wrong on rank {r}: max err {worst}"

    # a second, different job for the same scratch: 4 ranks, another shape
    world2, m2, n2, k2 = 4, 2048, 256, 1024
    rows2, kk2 = m2 // world2, k2 // world2
    a2 = torch.randn(m2, k2, device="cuda", dtype=torch.float16)
    b2 = torch.randn(k2, n2, device="cuda", dtype=torch.float16)
    exact2 = a2.float() @ b2.float()
    heaps2 = make_emulated_heaps(world2, 32 << 20)
    ops2 = [FusedGemmReduceScatter(heaps2[r], r, world2, m2, n2, k2, a2.dtype,
                                   a2[:, r * kk2:(r + 1) * kk2].contiguous(),
                                   b2[r * kk2:(r + 1) * kk2].contiguous()) for r in range(world2)]
    plans2 = (ctypes.c_void_p * world2)(*[op.plan._h for op in ops2])
    # first use on this device looks the scratch address up; do it before sampling memory
    N.check(lib.ddlb_tp_gemm_run_emulated_static(plans2, world2, 1, stream), "run_emulated_static")
    torch.cuda.synchronize()
    for r in range(world2):
        bad, worst = _within_tolerance(ops2[r].out, exact2[r * rows2:(r + 1) * rows2], k2)
        assert bad == 0, f"static entry point, other job, rank {r}: max err {worst}"

    # new entry point, two calls in a row on the shared static scratch, the other job in between
    free_before = torch.cuda.mem_get_info()[0]
    for epoch in (2, 3):
        for op in ops:
            op.out.fill_(float("nan"))
        N.check(lib.ddlb_tp_gemm_run_emulated_static(plans, world, epoch, stream),
                "run_emulated_static")
        torch.cuda.synchronize()
        for r in range(world):
            bad, worst = _within_tolerance(ops[r].out, expected[r], k)
            assert bad == 0, f"call with epoch {epoch}, rank {r}: {bad} elements off, max err {worst}"
        if epoch == 2:
            N.check(lib.ddlb_tp_gemm_run_emulated_static(plans2, world2, 2, stream),
                    "run_emulated_static")
            torch.cuda.synchronize()
    # nothing was taken from (or returned to) the device between the samples: the fills above
    # write existing tensors and the entry point itself allocates nothing
    assert torch.cuda.mem_get_info()[0] == free_before

    # more ranks than the fixed scratch holds: refused, with a message that says so
    too_many = (ctypes.c_void_p * 9)(*([ops[0].plan._h] * 9))
    rc = lib.ddlb_tp_gemm_run_emulated_static(too_many, 9, 4, stream)
    assert rc != 0
    message = lib.ddlb_last_error().decode()
    assert "exceeds" in message and "8" in message

    for op in ops + ops2:
        op.close()
    for h in heaps + heaps2:
        h.close()
