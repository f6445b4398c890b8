"""Roofline denominators and arithmetic.

``t_roof = max(flops_per_gpu / gemm_peak, comm_bytes_per_gpu / link_bw)``.
The GEMM peak is the MEASURED cuBLAS bf16 figure from ``MEASURED_PEAKS.json``
(burst, because every row here is a kernel timed in isolation); the link
bandwidth defaults to the nominal 900 GB/s per direction that BASELINE.json's
north star names, with the measured 770 GB/s peer copy reported alongside.
"""
from __future__ import annotations

import json
import os
from functools import lru_cache
from typing import Dict

FALLBACK_PEAKS = {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0}
NVLINK_NOMINAL_GBS = 900.0
NVLINK_MEASURED_GBS = 770.0


@lru_cache(maxsize=1)
def measured_peaks() -> Dict[str, float]:
    here = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    path = os.environ.get("DDLB_MEASURED_PEAKS", os.path.join(here, "MEASURED_PEAKS.json"))
    try:
        with open(path) as fh:
            data = json.load(fh)
        out = {k: float(data[k]) for k in FALLBACK_PEAKS if k in data}
        out["source"] = "measured"
        return {**FALLBACK_PEAKS, **out}
    except Exception:
        return {**FALLBACK_PEAKS, "source": "fallback"}


def gemm_peak_tflops(dtype: str) -> float:
    peak = measured_peaks()["bf16_tflops"]
    # tf32 tensor rate is half the 16-bit rate, and the fp32 path issues 3 MMAs per product
    return peak / 6.0 if dtype == "float32" else peak


def roofline_ms(flops_per_gpu: float, comm_bytes_per_gpu: float, dtype: str,
                link_gbs: float = NVLINK_NOMINAL_GBS) -> float:
    t_flop = flops_per_gpu / (gemm_peak_tflops(dtype) * 1e12)
    t_comm = comm_bytes_per_gpu / (link_gbs * 1e9)
    return max(t_flop, t_comm) * 1e3
