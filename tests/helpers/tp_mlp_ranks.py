"""Run under ``python -m ddlb_b200.launcher -np N``: the TP MLP block (AllGather->GEMM with the
activation fused, GEMM->ReduceScatter) on real ranks, several forward passes with NEW
activations each time (the input shard is overwritten at the top of every call, which the
closing handshake of the all-gather makes legal), against an fp32 reference.
Used by tests/test_multigpu.py."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.models import TPMLP  # noqa: E402


def main():
    comm = Communicator()
    d, r, dev = comm.world_size, comm.rank, comm.device
    heap = comm.symmetric_heap(512 << 20)
    tokens, hidden, ffn, dt = 1024 * d, 1024, 4096, torch.bfloat16
    rows, f_local = tokens // d, ffn // d
    torch.manual_seed(77)                                   # same weights on every rank
    w_up = torch.randn(hidden, ffn, device=dev, dtype=dt) * 0.05
    w_down = torch.randn(ffn, hidden, device=dev, dtype=dt) * 0.05
    mlp = TPMLP(heap, r, d, tokens, hidden, ffn, dt, w_up[:, r * f_local:(r + 1) * f_local].contiguous(),
                w_down[r * f_local:(r + 1) * f_local].contiguous(), activation="silu")
    torch.cuda.synchronize()
    heap.barrier()
    for it in range(6):
        torch.manual_seed(500 + it)
        x = torch.randn(tokens, hidden, device=dev, dtype=dt) * 0.5
        y = mlp(x[r * rows:(r + 1) * rows])                 # overwrites x_shard every call
        torch.cuda.synchronize()
        h = torch.nn.functional.silu(x.float() @ w_up.float()).to(dt)
        ref = (h.float() @ w_down.float())[r * rows:(r + 1) * rows]
        err = (y.float() - ref).abs().max().item()
        scale = ref.abs().max().item()
        assert err <= 3e-2 * scale, f"rank {r} pass {it}: max err {err} vs scale {scale}"
    assert mlp.gpu_launches == 12
    comm.barrier()
    print(f"tp-mlp ok rank {r}", flush=True)
    mlp.close()
    comm.shutdown()


if __name__ == "__main__":
    main()
