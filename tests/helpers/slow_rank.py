"""Run under ``python -m ddlb_b200.launcher -np N``: fused AG->GEMM (all transports), AG_after and
GEMM->RS with (a) one rank sleeping 200 ms before the first launch and before every third one,
and (b) inputs that change on every launch, checked against an fp32 product of the CURRENT
inputs. Used by tests/test_multigpu.py."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.ops.tp_gemm import FusedAllGatherGemm, FusedGemmReduceScatter  # noqa: E402


def check(out, ref, k, what):
    err = (out.float() - ref).abs()
    tol = 2e-2 * (k ** 0.5) + 1e-2 * ref.abs()
    bad = int((~(err <= tol)).sum())
    assert bad == 0, f"{what}: {bad}/{err.numel()} out of tolerance, max err {err.max().item()}"


def main():
    comm = Communicator()
    d, r, dev = comm.world_size, comm.rank, comm.device
    heap = comm.symmetric_heap(512 << 20)
    m, n, k, dt = 1024 * d, 512, 1024, torch.bfloat16
    rows, kk = m // d, k // d
    slow = d - 1                                           # the last rank drags its feet
    gen = torch.Generator(device=dev)
    b = torch.randn(k, n, device=dev, dtype=dt)            # same seed on every rank below
    torch.manual_seed(1234)
    b = torch.randn(k, n, device=dev, dtype=dt)
    transports = ["pull", "push"] + (["multimem"] if getattr(heap, "has_multicast", False) else [])
    cases = [("AG_before", t) for t in transports] + [("AG_after", "tma"), ("AG_after", "st128")]
    cases += [("AG_after", "multimem")] if "multimem" in transports else []
    for order, transport in cases:
        op = FusedAllGatherGemm(heap, r, d, m, n, k, dt, b, order=order, transport=transport)
        torch.cuda.synchronize()
        heap.barrier()
        for it in range(7):
            torch.manual_seed(100 + it)                    # every rank builds the same full A
            a_full = torch.randn(m, k, device=dev, dtype=dt)
            op.a_shard.copy_(a_full[r * rows:(r + 1) * rows])
            if r == slow and it % 3 == 0:
                torch.cuda.synchronize()
                time.sleep(0.2)
            out = op.run()
            torch.cuda.synchronize()
            check(out, a_full.float() @ b.float(), k, f"{order}/{transport} launch {it} rank {r}")
        op.close()
    rs_transports = ["push", "pull"] + (["multimem"] if "multimem" in transports else [])
    for transport in rs_transports:
        a_loc = torch.empty(m, kk, device=dev, dtype=dt)
        b_loc = torch.empty(kk, n, device=dev, dtype=dt)
        op = FusedGemmReduceScatter(heap, r, d, m, n, k, dt, a_loc, b_loc, transport=transport)
        torch.cuda.synchronize()
        heap.barrier()
        for it in range(7):
            torch.manual_seed(200 + it)
            a_full = torch.randn(m, k, device=dev, dtype=dt)
            b_full = torch.randn(k, n, device=dev, dtype=dt)
            a_loc.copy_(a_full[:, r * kk:(r + 1) * kk])
            b_loc.copy_(b_full[r * kk:(r + 1) * kk])
            if r == slow and it % 3 == 0:
                torch.cuda.synchronize()
                time.sleep(0.2)
            out = op.run()
            torch.cuda.synchronize()
            ref = (a_full.float() @ b_full.float())[r * rows:(r + 1) * rows]
            check(out, ref, k, f"RS/{transport} launch {it} rank {r}")
        op.close()
    torch.cuda.synchronize()
    comm.barrier()
    print(f"slow-rank ok rank {r} heap={type(heap).__name__} multicast={getattr(heap, 'has_multicast', False)}",
          flush=True)
    comm.shutdown()


if __name__ == "__main__":
    main()
