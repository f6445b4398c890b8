#!/usr/bin/env python3
"""Markdown tables for BASELINE.md section 4 from the records under profiles/r2/."""
import csv
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
P = os.path.join(ROOT, "profiles", "r2")


def load(name):
    with open(os.path.join(P, name)) as fh:
        return json.loads(fh.read().strip().splitlines()[-1])


def us(ms):
    return f"{ms * 1e3:.1f}"


def main():
    ns = (1, 2, 4, 8)
    ours = {n: load(f"bench_ours_n{n}.json") for n in ns}
    ref = {n: load(f"bench_reference_n{n}.json") for n in ns}
    out = []
    out.append("| tp_columnwise 65536x1024x1024 bf16, us per step (back to back) | N=1 | N=2 | N=4 | N=8 |")
    out.append("|---|---|---|---|---|")

    def row(label, fn):
        out.append(f"| {label} | " + " | ".join(fn(n) for n in ns) + " |")
    row("unmodified reference (`bench.py --impl reference`, its `empty_cache()` + `cudaMalloc` per step)",
        lambda n: us(ref[n]["ms_per_step"]))
    row("same, reference's default per-iteration-barrier protocol", lambda n: us(ref[n]["protocols"]["per_iteration_barrier"]))
    row("NCCL AllGather + cuBLAS, `empty_cache=false`, AG_before (the honest bar)",
        lambda n: us(ours[n]["baselines"]["nccl_allgather_cublas"]["ms_per_step"]))
    row("NCCL + cuBLAS, AG_after", lambda n: us(ours[n]["baselines"]["nccl_allgather_cublas_ag_after"]["ms_per_step"])
        if "nccl_allgather_cublas_ag_after" in ours[n]["baselines"] else "-")
    row("cuBLAS alone, full GEMM (`compute_only;size=unsharded`)",
        lambda n: us(ours[n]["baselines"]["cublas_only_unsharded"]["ms_per_step"]))
    row("**ours fused AG_before** (headline `value`)", lambda n: f"**{us(ours[n]['ms_per_step'])}**")
    row("ours, per-iteration-barrier protocol", lambda n: us(ours[n]["protocols"]["per_iteration_barrier"]))
    row("**ours fused AG_after**", lambda n: us(ours[n]["other_orders"]["ag_after"]["ms_per_step"])
        if ours[n].get("other_orders", {}).get("ag_after") else "-")
    row("ours AG_before vs NCCL+cuBLAS AG_before",
        lambda n: f"{ours[n]['baselines']['nccl_allgather_cublas']['ms_per_step'] / ours[n]['ms_per_step']:.2f}x")
    row("% of roofline (900 GB/s nominal / 770 GB/s measured link)",
        lambda n: f"{ours[n]['roofline']['pct_of_roofline']:.0f} / {ours[n]['roofline']['pct_of_roofline_measured_link']:.0f}")
    out.append("")
    out.append("| tp_rowwise 8192x1024x8192 fp16, us per step | N=1 | N=2 | N=4 | N=8 |")
    out.append("|---|---|---|---|---|")
    row("unmodified reference", lambda n: us(ref[n]["secondary"]["ms_per_step"]))
    row("cuBLAS + NCCL ReduceScatter, `empty_cache=false`",
        lambda n: us(ours[n]["baselines"]["cublas_nccl_reducescatter"]["ms_per_step"]))
    row("cuBLAS alone (what one rank computes)", lambda n: us(ours[n]["baselines"]["cublas_only_rowwise"]["ms_per_step"]))
    row("**ours fused** (transport)", lambda n: f"**{us(ours[n]['secondary']['ms_per_step'])}** ({ours[n]['secondary'].get('transport')})")
    row("ours vs cuBLAS+NCCL",
        lambda n: f"{ours[n]['baselines']['cublas_nccl_reducescatter']['ms_per_step'] / ours[n]['secondary']['ms_per_step']:.2f}x")
    row("% of roofline (nominal / measured link)",
        lambda n: f"{ours[n]['secondary']['pct_of_roofline']:.0f} / {ours[n]['secondary'].get('pct_of_roofline_measured_link', 0):.0f}")
    out.append("")
    out.append("| end to end (H2D of the input shard + step + D2H of the rank's result rows), ms per step | N=1 | N=2 | N=4 | N=8 |")
    out.append("|---|---|---|---|---|")
    row("bytes per step and rank, H2D / D2H (MiB)",
        lambda n: f"{ours[n]['e2e']['h2d_bytes_per_step'] / 2**20:.0f} / {ours[n]['e2e']['d2h_bytes_per_step'] / 2**20:.0f}")
    row("unmodified reference (serial, one stream)", lambda n: f"{ref[n]['e2e']['ms_per_step']:.2f}")
    row("ours, serial on one stream", lambda n: f"{ours[n]['e2e']['serial_ms_per_step']:.2f}")
    row("**ours, `PipelinedRunner`** (three streams)", lambda n: f"**{ours[n]['e2e']['ms_per_step']:.2f}**")
    row("ratio reference / ours", lambda n: f"{ref[n]['e2e']['ms_per_step'] / ours[n]['e2e']['ms_per_step']:.2f}x")
    out.append("")

    # sweep: fused vs library pair
    for d in (2, 4, 8):
        rows = list(csv.DictReader(open(os.path.join(P, f"sweep_d{d}.csv"))))
        by = {}
        for r in rows:
            by.setdefault((r["primitive"], int(r["m"]), int(r["n"]), int(r["k"])), {})[r["implementation"]] = float(r["ms"])
        wins = {"AG_before": [0, 0], "AG_after": [0, 0], "rowwise": [0, 0]}
        worst = []
        for (prim, m, n, k), v in by.items():
            if prim == "tp_columnwise":
                for order in ("AG_before", "AG_after"):
                    f, b = v.get(f"fused {order}"), v.get(f"nccl+cublas {order}")
                    if f and b:
                        wins[order][0 if f <= b else 1] += 1
                        worst.append((b / f, prim, order, m, n, k, f, b))
            else:
                f = next((x for kx, x in v.items() if kx.startswith("fused")), None)
                b = v.get("cublas+nccl reduce_scatter")
                if f and b:
                    wins["rowwise"][0 if f <= b else 1] += 1
                    worst.append((b / f, prim, "", m, n, k, f, b))
        out.append(f"Sweep at d={d} (`profiles/r2/sweep_d{d}.csv`, {len(by)} shapes incl. the Llama pair): fused faster / slower than "
                   f"the NCCL+cuBLAS pair - AG_before {wins['AG_before'][0]}/{wins['AG_before'][1]}, AG_after "
                   f"{wins['AG_after'][0]}/{wins['AG_after'][1]}, tp_rowwise {wins['rowwise'][0]}/{wins['rowwise'][1]}.")
        worst.sort()
        out.append("")
        out.append(f"| d={d}: five worst and five best shapes (library time / fused time) | m | n | k | fused us | library us | ratio |")
        out.append("|---|---|---|---|---|---|---|")
        for ratio, prim, order, m, n, k, f, b in worst[:5] + worst[-5:]:
            out.append(f"| {prim} {order} | {m} | {n} | {k} | {f * 1e3:.0f} | {b * 1e3:.0f} | {ratio:.2f}x |")
        out.append("")
        ll = [(key, v) for key, v in by.items() if key[1:] in ((16384, 28672, 8192), (16384, 8192, 28672))]
        if ll:
            out.append(f"| Llama-3-70B MLP shapes at d={d}, us | " + " | ".join(sorted({i for _, v in ll for i in v})) + " |")
            out.append("|---|" + "---|" * len(sorted({i for _, v in ll for i in v})))
            cols = sorted({i for _, v in ll for i in v})
            for key, v in ll:
                out.append(f"| {key[0]} {key[1]}x{key[2]}x{key[3]} | " + " | ".join(f"{v[c] * 1e3:.0f}" if c in v else "-" for c in cols) + " |")
            out.append("")
    print("\n".join(out))


if __name__ == "__main__":
    main()
