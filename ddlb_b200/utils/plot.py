"""Headless bar chart (replaces the reference's ``plt.show()``,
``ddlb/benchmark.py:391-425``). PNG via matplotlib/Agg when importable, else a
small hand-written SVG so the feature never depends on an optional package."""
from __future__ import annotations

import html
import math
import os
from typing import List


def _svg(labels: List[str], values: List[float], errors: List[float], title: str,
         ylabel: str) -> str:
    w, h, left, bottom, top = 960, 540, 80, 170, 50
    plot_w, plot_h = w - left - 30, h - bottom - top
    finite = [v + (e if math.isfinite(e) else 0) for v, e in zip(values, errors) if math.isfinite(v)]
    vmax = max(finite) if finite else 1.0
    vmax = vmax * 1.1 if vmax > 0 else 1.0
    n = max(len(values), 1)
    slot = plot_w / n
    bar_w = slot * 0.6
    out = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{w}" height="{h}" '
           f'font-family="sans-serif" font-size="12">',
           f'<rect width="{w}" height="{h}" fill="white"/>',
           f'<text x="{w / 2}" y="24" text-anchor="middle" font-size="16">{html.escape(title)}</text>',
           f'<text x="16" y="{top + plot_h / 2}" transform="rotate(-90 16 {top + plot_h / 2})" '
           f'text-anchor="middle">{html.escape(ylabel)}</text>',
           f'<line x1="{left}" y1="{top + plot_h}" x2="{left + plot_w}" y2="{top + plot_h}" stroke="black"/>',
           f'<line x1="{left}" y1="{top}" x2="{left}" y2="{top + plot_h}" stroke="black"/>']
    for i in range(5):
        yv = vmax * i / 4
        y = top + plot_h - plot_h * i / 4
        out.append(f'<text x="{left - 6}" y="{y + 4}" text-anchor="end">{yv:.3g}</text>')
        out.append(f'<line x1="{left}" y1="{y}" x2="{left + plot_w}" y2="{y}" stroke="#ddd"/>')
    for i, (lab, v, e) in enumerate(zip(labels, values, errors)):
        if not math.isfinite(v):
            continue
        x = left + slot * i + (slot - bar_w) / 2
        bh = plot_h * v / vmax
        y = top + plot_h - bh
        out.append(f'<rect x="{x:.1f}" y="{y:.1f}" width="{bar_w:.1f}" height="{bh:.1f}" fill="#4878a8"/>')
        if math.isfinite(e) and e > 0:
            cx, eh = x + bar_w / 2, plot_h * e / vmax
            out.append(f'<line x1="{cx:.1f}" y1="{y - eh:.1f}" x2="{cx:.1f}" y2="{y + eh:.1f}" stroke="black"/>')
        tx, ty = x + bar_w / 2, top + plot_h + 12
        out.append(f'<text x="{tx:.1f}" y="{ty}" transform="rotate(45 {tx:.1f} {ty})">'
                   f'{html.escape(str(lab))}</text>')
    out.append("</svg>")
    return "\n".join(out)


def bar_chart(labels: List[str], values: List[float], errors: List[float], title: str,
              ylabel: str, path_stem: str) -> str:
    directory = os.path.dirname(path_stem)
    if directory:
        os.makedirs(directory, exist_ok=True)
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
        fig = plt.figure(figsize=(12, 6))
        plt.bar(labels, values, yerr=errors, capsize=5)
        plt.title(title)
        plt.ylabel(ylabel)
        plt.xticks(rotation=45, ha="right")
        plt.tight_layout()
        path = path_stem + ".png"
        fig.savefig(path)
        plt.close(fig)
        return path
    except ImportError:
        path = path_stem + ".svg"
        with open(path, "w") as fh:
            fh.write(_svg(labels, [float(v) for v in values], [float(e) for e in errors],
                          title, ylabel))
        return path
