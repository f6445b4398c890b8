"""Device trace of a fused launch (first run on a B200 in round 2: passed)."""
import pytest

pytestmark = pytest.mark.gpu


def test_trace_single_gpu_plain_rs():
    import torch
    from ddlb_b200.ops.tp_gemm import FusedGemmReduceScatter
    from ddlb_b200.utils.trace import decode, summarize
    a = torch.randn(2048, 512, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(512, 512, device="cuda", dtype=torch.bfloat16)
    op = FusedGemmReduceScatter(None, 0, 1, 2048, 512, 512, a.dtype, a, b, cta_group=2,
                                block_n=256, trace=True)
    out = op.run()
    torch.cuda.synchronize()
    torch.testing.assert_close(out.float(), a.float() @ b.float(), rtol=2e-2, atol=1.0)
    events = decode(op.trace_buf)
    tags = {e["tag"] for e in events}
    assert {"start", "tile_begin", "tile_issued", "acc_ready", "tile_stored", "end"} <= tags
    assert sum(1 for e in events if e["tag"] == "tile_stored") == (2048 // 128) * 2   # per CTA tile
    print(summarize(events))
