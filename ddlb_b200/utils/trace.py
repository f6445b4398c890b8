"""Opt-in device trace of one fused launch (timeline without a profiler).

``ncu`` cannot replay kernels that wait for other ranks, so the fused kernels can record
``globaltimer`` stamps at phase boundaries themselves (``csrc/tp_gemm_sm100.cuh``, ``Tracer``;
compiled into a separate kernel instantiation, production kernels carry no trace code).
Available for fp16/bf16, cta_group=2, block_n=256, B ``[k,n]``.

    op = FusedGemmReduceScatter(..., trace=True)
    op.run(); torch.cuda.synchronize()
    print(summarize(decode(op.trace_buf)))
"""
from __future__ import annotations

from typing import Dict, List

from .. import _native as N

ROLES = {0: "producer", 1: "mma", 2: "epilogue", 3: "comm"}
TAGS = {1: "start", 2: "tile_begin", 3: "flag_waited", 4: "tile_issued", 5: "acc_ready",
        6: "unit_ready", 7: "tile_stored", 8: "unit_done", 9: "end"}


def alloc_trace_buffer(device, max_ctas: int = 160):
    """Zeroed buffer for ``max_ctas`` CTAs (16-byte records)."""
    import torch
    return torch.zeros(max_ctas * N.TRACE_ROLES * N.TRACE_SLOTS, 4, dtype=torch.int32, device=device)


def decode(buf) -> List[Dict]:
    """Records with a non-zero tag, as dicts with ``cta, role, tag, arg, t_ns`` (t relative
    to the earliest stamp)."""
    raw = buf.cpu().numpy().astype("uint32")
    events = []
    per_cta = N.TRACE_ROLES * N.TRACE_SLOTS
    for idx in raw[:, 0].nonzero()[0]:
        tag, arg, lo, hi = (int(x) for x in raw[idx])
        events.append({"cta": int(idx // per_cta), "role": ROLES.get(int(idx % per_cta) // N.TRACE_SLOTS),
                       "tag": TAGS.get(tag, str(tag)), "arg": arg, "t_ns": (hi << 32) | lo})
    if events:
        t0 = min(e["t_ns"] for e in events)
        for e in events:
            e["t_ns"] -= t0
    events.sort(key=lambda e: e["t_ns"])
    return events


def summarize(events: List[Dict]) -> str:
    """Per tag: count, first, median and last time in microseconds."""
    by_tag: Dict[str, List[int]] = {}
    for e in events:
        by_tag.setdefault(f'{e["role"]}:{e["tag"]}', []).append(e["t_ns"])
    lines = [f'{"event":28s} {"n":>6s} {"first_us":>10s} {"median_us":>10s} {"last_us":>10s}']
    for key, ts in sorted(by_tag.items(), key=lambda kv: min(kv[1])):
        ts.sort()
        lines.append(f"{key:28s} {len(ts):6d} {ts[0] / 1e3:10.2f} {ts[len(ts) // 2] / 1e3:10.2f} "
                     f"{ts[-1] / 1e3:10.2f}")
    return "\n".join(lines)
