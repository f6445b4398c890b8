"""TP MLP block (AllGather->GEMM with the activation fused into its epilogue, GEMM->ReduceScatter)
on one GPU: world=1 through the fused kernels, and two emulated ranks, against a PyTorch fp32
reference."""
import pytest

pytestmark = pytest.mark.gpu


def _reference(x, w_up, w_down, act="silu"):
    import torch
    pre = x.float() @ w_up.float()                       # the kernel activates the fp32 accumulator
    fn = {"silu": torch.nn.functional.silu, "relu": torch.relu,
          "gelu_tanh": lambda t: torch.nn.functional.gelu(t, approximate="tanh")}[act]
    h = fn(pre).to(x.dtype)
    return h.float() @ w_down.float()


@pytest.mark.parametrize("act", ["silu", "relu", "gelu_tanh"])
def test_tp_mlp_single_rank(act):
    import torch
    from ddlb_b200.models import TPMLP
    torch.manual_seed(5)
    tokens, hidden, ffn, dt = 512, 256, 1024, torch.bfloat16
    x = torch.randn(tokens, hidden, device="cuda", dtype=dt) * 0.5
    w_up = torch.randn(hidden, ffn, device="cuda", dtype=dt) * 0.1
    w_down = torch.randn(ffn, hidden, device="cuda", dtype=dt) * 0.1
    mlp = TPMLP(None, 0, 1, tokens, hidden, ffn, dt, w_up, w_down, activation=act)
    y = mlp(x)
    torch.cuda.synchronize()
    ref = _reference(x, w_up, w_down, act)
    torch.testing.assert_close(y.float(), ref, rtol=3e-2, atol=3e-2 * ref.abs().max().item())
    assert mlp.gpu_launches == 2
    mlp.close()


def test_tp_mlp_two_emulated_ranks():
    import torch
    from ddlb_b200.models import TPMLP
    from ddlb_b200.ops.emulation import make_emulated_heaps, run_emulated
    torch.manual_seed(6)
    world, tokens, hidden, ffn, dt = 2, 1024, 256, 1024, torch.bfloat16
    rows, f_local = tokens // world, ffn // world
    x = torch.randn(tokens, hidden, device="cuda", dtype=dt) * 0.5
    w_up = torch.randn(hidden, ffn, device="cuda", dtype=dt) * 0.1
    w_down = torch.randn(ffn, hidden, device="cuda", dtype=dt) * 0.1
    heaps = make_emulated_heaps(world, 64 << 20)
    mlps = [TPMLP(heaps[r], r, world, tokens, hidden, ffn, dt,
                  w_up[:, r * f_local:(r + 1) * f_local].contiguous(),
                  w_down[r * f_local:(r + 1) * f_local].contiguous()) for r in range(world)]
    for r, m in enumerate(mlps):
        m.x_shard.copy_(x[r * rows:(r + 1) * rows])
    torch.cuda.synchronize()
    # the ranks of one emulated launch must run together: drive the two phases by hand
    run_emulated([m.up.plan for m in mlps], 1)
    run_emulated([m.down.plan for m in mlps], 1)      # H went straight from one kernel to the next
    torch.cuda.synchronize()
    ref = _reference(x, w_up, w_down)
    for r, m in enumerate(mlps):
        got = m.down.out.float()
        want = ref[r * rows:(r + 1) * rows]
        torch.testing.assert_close(got, want, rtol=3e-2, atol=3e-2 * ref.abs().max().item())
    for m in mlps:
        m.close()
    for h in heaps:
        h.close()
