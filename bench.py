#!/usr/bin/env python3
"""Headline benchmark (driver contract).

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference]

N > 1 is launched by the driver with ``python -m torch.distributed.run --nproc-per-node N
... bench.py --gpus N ...`` (one rank per GPU). Rank 0 prints ONE JSON line.

Metric (BASELINE.json): tensor-parallel AllGather->GEMM throughput, ``2*m*n*k / t``
TFLOP/s as in the reference's ``Throughput (TFLOPS)`` column
(``ddlb/benchmark.py:209-214``), on the headline config ``tp_columnwise m=65536 n=1024
k=1024 bf16``; a "step" is one ``run()`` of the primitive (one fused kernel launch per
rank for our arm; ncclAllGather + cuBLAS for the reference arm). The global problem
is fixed as N grows ("strong" scaling). ``secondary`` reports the GEMM->ReduceScatter
headline config (``tp_rowwise m=8192 n=1024 k=8192 fp16``) measured the same way.

``--impl reference`` runs the UNMODIFIED reference from ``baseline/_ref`` through its
own ``ddlb.cli.run_benchmark`` (``pytorch;backend=nccl;order=AG_before``, CUDA-event
timing) and prints the same JSON shape with ``"impl": "reference"``.

Both arms report ``value`` with the same protocol (one CUDA-event pair around K back-to-back
steps), list the reference's default per-iteration-barrier protocol under ``protocols``,
and carry an ``e2e`` block: every step copies the rank's input shard from pinned host
memory, runs the primitive and reads the rank's rows of the result back. Our arm adds a
``baselines`` block (NCCL + cuBLAS without the reference's in-loop ``empty_cache()``, and
cuBLAS alone) measured in the same process with the same protocols.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HEADLINE = {"primitive": "tp_columnwise", "m": 65536, "n": 1024, "k": 1024, "dtype": "bfloat16"}
SECONDARY = {"primitive": "tp_rowwise", "m": 8192, "n": 1024, "k": 8192, "dtype": "float16"}


# ----------------------------------------------------------------------------- clocks
class ClockSampler:
    """Sample SM clock / throttle reasons of one GPU through NVML while the timed
    region runs (a thread; ``torch.cuda.synchronize`` releases the GIL)."""

    REASONS = {
        0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
        0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
        0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting",
    }

    def __init__(self, index: int, period_s: float = 0.004):
        self.index, self.period = index, period_s
        self.samples, self.reason_bits = [], 0
        self.max_mhz = None
        self._stop = threading.Event()
        self._thread = None
        self._nvml = None
        self._handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._handle, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._nvml = None

    def sample(self) -> None:
        if self._nvml is None:
            return
        try:
            n = self._nvml
            self.samples.append(n.nvmlDeviceGetClockInfo(self._handle, n.NVML_CLOCK_SM))
            self.reason_bits |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self._handle))
        except Exception:
            pass

    def _loop(self) -> None:
        while not self._stop.is_set():
            self.sample()
            time.sleep(self.period)

    def start(self) -> None:
        self.samples, self.reason_bits = [], 0
        self._stop.clear()
        self.sample()
        if self._nvml is not None:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join()
            self._thread = None
        self.sample()
        reasons = [name for bit, name in self.REASONS.items()
                   if self.reason_bits & bit and name != "gpu_idle"]
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": reasons, "samples": len(self.samples)}


def _physical_gpu_index(local_rank: int) -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            pass
    return local_rank


# -------------------------------------------------------------------------- our arm
def _timed_steps(impl, comm, steps: int):
    """EXACTLY ``steps`` runs between barrier+sync brackets, CUDA events; ms per step."""
    import torch
    comm.barrier()
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    out = None
    for _ in range(steps):
        out = impl.run()
    end.record()
    torch.cuda.synchronize()
    comm.barrier()
    return start.elapsed_time(end) / steps, out


def _timed_per_iteration(impl, comm, steps: int):
    """The reference's default protocol (``ddlb/benchmark.py:127-144``): a cross-rank barrier,
    then one event pair per iteration; mean over iterations of the max over ranks. The barrier
    is the stream-ordered device barrier for every implementation in this arm."""
    import torch
    heap = comm.symmetric_heap() if comm.world_size > 1 else None
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    comm.barrier()
    for i in range(steps):
        if heap is not None:
            heap.barrier()
        starts[i].record()
        impl.run()
        ends[i].record()
    torch.cuda.synchronize()
    times = comm.all_reduce_max([s.elapsed_time(e) for s, e in zip(starts, ends)])
    return sum(times) / len(times)


def _result_rows(comm, out, m: int):
    """The rows of the replicated result that belong to this rank's input shard: what a
    sequence-parallel consumer of this rank reads back."""
    rows = m // comm.world_size
    return out[comm.rank * rows:(comm.rank + 1) * rows]


def _e2e_serial(impl, comm, cfg: dict, steps: int, warmup: int):
    """End to end on one stream, every step: H2D of the input shard from pinned memory, run,
    D2H of this rank's result rows."""
    import torch
    shard_dev = impl.A                                    # device tensor the kernel reads
    host_in = torch.empty(shard_dev.shape, dtype=shard_dev.dtype).pin_memory()
    host_in.copy_(shard_dev)
    probe = _result_rows(comm, impl.run(), cfg["m"])
    host_out = torch.empty(probe.shape, dtype=probe.dtype).pin_memory()
    h2d = host_in.numel() * host_in.element_size()
    d2h = host_out.numel() * host_out.element_size()

    def step():
        shard_dev.copy_(host_in, non_blocking=True)
        out = impl.run()
        host_out.copy_(_result_rows(comm, out, cfg["m"]), non_blocking=True)

    for _ in range(warmup):
        step()
    comm.barrier()
    torch.cuda.synchronize()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(steps):
        step()
    end.record()
    torch.cuda.synchronize()
    comm.barrier()
    return start.elapsed_time(end) / steps, h2d, d2h


def _e2e_pipelined(cfg: dict, comm, steps: int, warmup: int):
    """Same bytes per step through ``ddlb_b200.utils.pipeline.PipelinedRunner`` (two rotating
    instances of the primitive): the H2D copy of step i+1 and the D2H read of step i-1
    overlap the kernel of step i (PCIe is full duplex)."""
    import torch
    from ddlb_b200.primitives.registry import load_impl_class
    from ddlb_b200.utils.pipeline import PipelinedRunner
    impl_cls = load_impl_class(cfg["primitive"], "fused")
    runner = PipelinedRunner(lambda: impl_cls(m=cfg["m"], n=cfg["n"], k=cfg["k"],
                                              dtype=cfg["dtype"]), depth=2)
    try:
        rows = cfg["m"] // comm.world_size
        sl = slice(comm.rank * rows, (comm.rank + 1) * rows)
        host_in = torch.empty(runner.input_shape, dtype=runner.input_dtype).pin_memory()
        host_in.copy_(runner.impls[0].A)
        host_out = torch.empty(rows, cfg["n"], dtype=runner.input_dtype).pin_memory()
        h2d = host_in.numel() * host_in.element_size()
        d2h = host_out.numel() * host_out.element_size()
        for _ in range(max(warmup, 3)):
            runner.step(host_in, host_out, sl)
        comm.barrier()
        torch.cuda.synchronize()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record()
        out = None
        for _ in range(steps):
            out = runner.step(host_in, host_out, sl)
        runner.drain()
        end.record()
        torch.cuda.synchronize()
        comm.barrier()
        ms = start.elapsed_time(end) / steps
        last = runner.impls[(runner._step - 1) % runner.depth]
        last.validate(out, strict=True)
        # the bytes that reached the host are the rank's rows of that result
        ok = bool(torch.equal(host_out, out[sl].cpu()))
        return ms, h2d, d2h, ok, runner.gpu_launches
    finally:
        runner.close()


def _measure(cfg: dict, impl_name: str, options: dict, steps: int, warmup: int, sampler=None,
             per_iteration: bool = True, strict: bool = True):
    """Kernel-only measurement of one implementation: contract protocol (back to back) and
    the reference's per-iteration protocol; validation on a poisoned extra launch."""
    import torch
    from ddlb_b200.primitives.registry import load_impl_class
    impl_cls = load_impl_class(cfg["primitive"], impl_name)
    impl = impl_cls(m=cfg["m"], n=cfg["n"], k=cfg["k"], dtype=cfg["dtype"], **options)
    comm = impl.communicator
    for _ in range(max(warmup, 3)):
        impl.run()
    torch.cuda.synchronize()
    launches0 = getattr(impl, "gpu_launches", 0)
    if sampler is not None:
        sampler.start()
    ms, out = _timed_steps(impl, comm, steps)
    clocks = sampler.stop() if sampler is not None else None
    launches = getattr(impl, "gpu_launches", 0) - launches0
    ms = max(comm.all_reduce_max([ms]))
    res = {"ms_per_step": ms, "gpu_launches": launches, "clocks": clocks,
           "flops_per_gpu": impl.flops_per_gpu(), "comm_bytes_per_gpu": impl.comm_bytes_per_gpu()}
    if per_iteration:
        res["ms_per_step_per_iteration_barrier"] = _timed_per_iteration(impl, comm, steps)
    op = getattr(impl, "op", None)
    if op is not None:
        res["plan"] = dict(op.plan.info) if getattr(op, "plan", None) else {}
        res["transport"] = getattr(op, "transport", None)
    valid = True
    try:
        torch.cuda.synchronize()
        impl.poison()
        out = impl.run()
        torch.cuda.synchronize()
        impl.validate(out, strict=strict and op is not None)
    except Exception as exc:
        valid = False
        print(f"[bench] validation failed ({impl_name} {options}): {str(exc).splitlines()[0]}",
              file=sys.stderr)
    ok = min(-v for v in comm.all_reduce_max([-(1.0 if valid else 0.0)])) > 0.5
    res["valid"] = ok
    return res, impl


def run_ours(args) -> dict:
    os.environ.setdefault("DDLB_SYMM_HEAP_BYTES", str(3 << 30))
    import torch
    from ddlb_b200.communicator import Communicator
    from ddlb_b200.utils.roofline import measured_peaks, roofline_ms, NVLINK_MEASURED_GBS

    comm = Communicator()
    if comm.world_size != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE is {comm.world_size}: launch with "
                         f"torch.distributed.run --nproc-per-node {args.gpus}")
    sampler = ClockSampler(_physical_gpu_index(comm.local_rank))

    def tflops(cfg, ms):
        return 2.0 * cfg["m"] * cfg["n"] * cfg["k"] / (ms * 1e9)

    # ---- headline: fused AllGather->GEMM, order AG_before (BASELINE.json), kernel only
    head, impl = _measure(HEADLINE, "fused", {"order": "AG_before"}, args.steps, args.warmup, sampler)
    e_ser_ms, h2d, d2h = _e2e_serial(impl, comm, HEADLINE, args.steps, max(args.warmup, 3))
    e_ser_ms = max(comm.all_reduce_max([e_ser_ms]))
    heap_kind = type(comm.symmetric_heap()).__name__ if comm.world_size > 1 else None
    multicast = bool(getattr(comm.symmetric_heap(), "has_multicast", False)) if comm.world_size > 1 else False
    impl.close()
    # ---- end to end through the public pipelined runner (H2D + kernel + D2H every step)
    e2e = {"serial_ms_per_step": e_ser_ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}
    try:
        p_ms, h2d, d2h, bytes_ok, p_launches = _e2e_pipelined(HEADLINE, comm, args.steps, args.warmup)
        p_ms = max(comm.all_reduce_max([p_ms]))
        e2e.update({"ms_per_step": p_ms, "mode": "PipelinedRunner(depth=2): H2D, kernel, D2H on three streams",
                    "valid": bytes_ok})
    except Exception as exc:                       # keep the serial number, say why
        print(f"[bench] pipelined e2e unavailable: {str(exc).splitlines()[0][:200]}", file=sys.stderr)
        e2e.update({"ms_per_step": e_ser_ms, "mode": "serial (one stream)", "valid": head["valid"]})
    e2e["value"] = tflops(HEADLINE, e2e["ms_per_step"])
    e2e["unit"] = "TFLOP/s (2*m*n*k/t, whole job)"
    e2e["result_read_back"] = "this rank's m/N rows of C (the rows of its own input shard)"

    # ---- the same primitive in the other order, and the second primitive
    extra = {}
    if comm.world_size > 1:
        after, impl2 = _measure(HEADLINE, "fused", {"order": "AG_after"}, args.steps, args.warmup)
        impl2.close()
        extra["ag_after"] = {"ms_per_step": after["ms_per_step"], "value": tflops(HEADLINE, after["ms_per_step"]),
                             "valid": after["valid"], "transport": after.get("transport"),
                             "ms_per_step_per_iteration_barrier": after.get("ms_per_step_per_iteration_barrier")}
    sec, impl3 = _measure(SECONDARY, "fused", {}, args.steps, args.warmup)
    impl3.close()

    # ---- side-by-side library baselines at the same N, same process, same protocols: the
    # reference's algorithm WITHOUT its in-loop torch.cuda.empty_cache(), and the compute ceiling
    baselines = {}
    if os.environ.get("DDLB_BENCH_BASELINES", "1") != "0":
        for key, cfg, name, opts in (
                ("nccl_allgather_cublas", HEADLINE, "pytorch", {"backend": "nccl", "order": "AG_before", "empty_cache": False}),
                ("nccl_allgather_cublas_ag_after", HEADLINE, "pytorch", {"backend": "nccl", "order": "AG_after", "empty_cache": False}),
                ("cublas_only_unsharded", HEADLINE, "compute_only", {"size": "unsharded"}),
                ("cublas_nccl_reducescatter", SECONDARY, "pytorch", {"backend": "nccl", "empty_cache": False}),
                ("cublas_only_rowwise", SECONDARY, "compute_only", {"size": "sharded"})):
            if comm.world_size == 1 and key == "nccl_allgather_cublas_ag_after":
                continue
            try:
                r, b_impl = _measure(cfg, name, opts, args.steps, args.warmup, strict=False)
                b_impl.close()
                baselines[key] = {"ms_per_step": r["ms_per_step"], "value": tflops(cfg, r["ms_per_step"]),
                                  "ms_per_step_per_iteration_barrier": r.get("ms_per_step_per_iteration_barrier"),
                                  "valid": r["valid"]}
            except Exception as exc:
                baselines[key] = {"unavailable": f"{type(exc).__name__}: {str(exc).splitlines()[0][:160]}"}

    roof = roofline_ms(head["flops_per_gpu"], head["comm_bytes_per_gpu"], HEADLINE["dtype"])
    roof_meas = roofline_ms(head["flops_per_gpu"], head["comm_bytes_per_gpu"], HEADLINE["dtype"],
                            link_gbs=NVLINK_MEASURED_GBS)
    sec_roof = roofline_ms(sec["flops_per_gpu"], sec["comm_bytes_per_gpu"], SECONDARY["dtype"])
    sec_roof_meas = roofline_ms(sec["flops_per_gpu"], sec["comm_bytes_per_gpu"], SECONDARY["dtype"],
                                link_gbs=NVLINK_MEASURED_GBS)
    peaks = measured_peaks()
    out = {
        "metric": "tp_columnwise_tflops", "value": tflops(HEADLINE, head["ms_per_step"]),
        "unit": "TFLOP/s (2*m*n*k/t, whole job)", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "impl": "ours", "valid": head["valid"],
        "config": _common_config(args.gpus),
        "implementation": {"name": "fused (one sm_100a kernel per rank, no NCCL)", "kernel": head.get("plan"),
                           "transport": head.get("transport"), "symmetric_heap": heap_kind,
                           "nvls_multicast_window": multicast,
                           "nccl": "used by the `baselines` block only, after every fused measurement"},
        "protocols": {"back_to_back (value)": head["ms_per_step"],
                      "per_iteration_barrier": head.get("ms_per_step_per_iteration_barrier")},
        "clocks": head["clocks"],
        "gpu_launches": head["gpu_launches"],
        "e2e": e2e,
        "roofline": {"t_roof_ms": roof, "pct_of_roofline": 100.0 * roof / head["ms_per_step"],
                     "t_roof_ms_measured_link": roof_meas,
                     "pct_of_roofline_measured_link": 100.0 * roof_meas / head["ms_per_step"],
                     "gemm_peak_tflops": peaks["bf16_tflops"], "peaks_source": peaks["source"]},
        "secondary": {"metric": "tp_rowwise_tflops",
                      "config": "tp_rowwise GEMM->ReduceScatter m=8192 n=1024 k=8192 fp16",
                      "value": tflops(SECONDARY, sec["ms_per_step"]),
                      "ms_per_step": sec["ms_per_step"], "valid": sec["valid"],
                      "ms_per_step_per_iteration_barrier": sec.get("ms_per_step_per_iteration_barrier"),
                      "gpu_launches": sec["gpu_launches"], "transport": sec.get("transport"),
                      "pct_of_roofline": 100.0 * sec_roof / sec["ms_per_step"],
                      "pct_of_roofline_measured_link": 100.0 * sec_roof_meas / sec["ms_per_step"]},
        "other_orders": extra,
        "baselines": baselines,
    }
    comm.shutdown()
    return out if comm.rank == 0 else {}


def _common_config(gpus: int) -> dict:
    """Identical in both arms: what is measured, not who measures it."""
    return {"model": "tp_columnwise AllGather->GEMM m=65536 n=1024 k=1024",
            "global_batch": HEADLINE["m"], "seq_len": HEADLINE["k"],
            "parallelism": f"tp{gpus}", "order": "AG_before",
            "timing": "CUDA events around K back-to-back steps between barrier+synchronize brackets, max over ranks",
            "l2": "working set (A 128 MiB + C 128 MiB) exceeds the 126 MB L2; no flush"}


# -------------------------------------------------------------------- reference arm
def _ensure_reference():
    ref = os.path.join(ROOT, "baseline", "_ref")
    if os.path.isdir(os.path.join(ref, "ddlb")):
        return ref, None
    src = "/root/reference"
    if not os.path.isdir(src):
        return None, "baseline/_ref is missing and /root/reference is not present on this box"
    import shutil
    import tempfile
    tmp = tempfile.mkdtemp(prefix="ddlb_ref_")
    copy = os.path.join(tmp, "reference")
    shutil.copytree(src, copy)
    cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
           "--no-deps", "--find-links", "/opt/wheelhouse", "--target", ref, copy]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if proc.returncode != 0:
        return None, "pip install of the reference failed: " + proc.stderr.strip().splitlines()[-1][:200]
    return ref, None


def _reference_case(ddlb_cli, cfg: dict, steps: int, warmup: int, csv_path: str, rank: int,
                    barrier_each_iteration: bool):
    import csv
    impl_opts = ({"backend": "nccl", "order": "AG_before"} if cfg["primitive"] == "tp_columnwise"
                 else {"backend": "nccl"})
    config = {"benchmark": {
        "primitive": cfg["primitive"], "m": cfg["m"], "n": cfg["n"], "k": cfg["k"],
        "dtype": cfg["dtype"], "validate": False, "num_iterations": steps,
        "num_warmups": warmup, "output_csv": csv_path,
        "time_measurement_backend": "cuda_event",
        "barrier_at_each_iteration": barrier_each_iteration,
        "implementations": {"pytorch": [impl_opts]}}}
    ddlb_cli.run_benchmark(config)
    if rank != 0:
        return None
    with open(csv_path) as fh:
        rows = list(csv.DictReader(fh))
    row = rows[-1]
    return {"ms_per_step": float(row["mean_time (ms)"]), "tflops": float(row["Throughput (TFLOPS)"]),
            "label": row["implementation"]}


def _reference_e2e(cfg: dict, steps: int, warmup: int, rank: int, world: int):
    """End to end through the reference's own public class, unmodified: every step copies this
    rank's input shard from pinned host memory into ``impl.A``, calls
    ``PyTorchTPColumnwise.run()`` (``ddlb/primitives/TPColumnwise/pytorch.py:85-104``: its
    ``empty_cache`` + ncclAllGather + cuBLAS) and reads this rank's rows of the result back -
    the same bytes per step as our arm. One stream: the reference has no overlapped runner."""
    import torch
    from ddlb.primitives.TPColumnwise import PyTorchTPColumnwise
    impl = PyTorchTPColumnwise(cfg["m"], cfg["n"], cfg["k"], dtype=cfg["dtype"], backend="nccl",
                               order="AG_before")
    rows = cfg["m"] // world
    sl = slice(rank * rows, (rank + 1) * rows)
    host_in = torch.empty(impl.A.shape, dtype=impl.A.dtype).pin_memory()
    host_in.copy_(impl.A)
    host_out = torch.empty(rows, cfg["n"], dtype=impl.A.dtype).pin_memory()
    h2d = host_in.numel() * host_in.element_size()
    d2h = host_out.numel() * host_out.element_size()

    def step():
        impl.A.copy_(host_in, non_blocking=True)
        out = impl.run()
        host_out.copy_(out[sl], non_blocking=True)

    for _ in range(max(warmup, 3)):
        step()
    torch.cuda.synchronize()
    impl.communicator.barrier()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(steps):
        step()
    end.record()
    torch.cuda.synchronize()
    ms = start.elapsed_time(end) / steps
    import torch.distributed as dist
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=impl.A.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    del impl                                           # its __del__ destroys the process group
    return {"ms_per_step": ms, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}


def _arm_watchdog(seconds: float, payload: dict) -> None:
    """The reference's parent blocks forever on ``result_queue.get()`` when its child dies
    before posting a row (``ddlb/benchmark.py:369``). Never let that hang the driver:
    after ``seconds`` report the arm as unavailable, stop our own children and exit 0."""
    def fire():
        import multiprocessing as mp
        try:
            for child in mp.active_children():
                child.terminate()
        except Exception:
            pass
        if int(os.environ.get("RANK", "0")) == 0:
            os.write(1 if not hasattr(_arm_watchdog, "fd") else _arm_watchdog.fd,
                     (json.dumps(payload) + "\n").encode())
        os._exit(0)
    timer = threading.Timer(seconds, fire)
    timer.daemon = True
    timer.start()


def run_reference(args) -> dict:
    try:
        import torch
        if not torch.cuda.is_available():
            return {"impl": "reference", "unavailable": "no CUDA device: the reference's "
                    "Communicator requires one (ddlb/communicator.py:56-60)"}
    except Exception as exc:
        return {"impl": "reference", "unavailable": f"torch import failed: {exc}"}
    ref, why = _ensure_reference()
    if ref is None:
        return {"impl": "reference", "unavailable": why}
    _arm_watchdog(float(os.environ.get("DDLB_REFERENCE_TIMEOUT", "600")),
                  {"impl": "reference", "unavailable": "reference run did not finish in time "
                   "(its parent hangs if a child dies before posting a row)"})
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        return {"impl": "reference", "unavailable": f"--gpus {args.gpus} but WORLD_SIZE={world}"}
    # the reference reads Open MPI / SLURM variables only (ddlb/envs.py:50-67)
    os.environ.update({"OMPI_COMM_WORLD_RANK": str(rank), "OMPI_COMM_WORLD_SIZE": str(world),
                       "OMPI_COMM_WORLD_LOCAL_RANK": str(local_rank),
                       "OMPI_COMM_WORLD_LOCAL_SIZE": os.environ.get("LOCAL_WORLD_SIZE", str(world)),
                       "DDLB_MASTER_ADDR": "127.0.0.1",
                       "DDLB_MASTER_PORT": str(int(os.environ.get("MASTER_PORT", "29500")) + 307)})
    # torchrun marks its workers with TORCHELASTIC_USE_AGENT_STORE=True; inherited by the
    # reference's children it turns EVERY rank of their tcp:// rendezvous into a client of a
    # store nobody hosts on DDLB_MASTER_PORT (torch.distributed.rendezvous._create_c10d_store),
    # and the run waits for the connect timeout. The reference expects an mpirun-style
    # environment, so give it one.
    for key in [k for k in os.environ if k.startswith("TORCHELASTIC_")]:
        os.environ.pop(key, None)
    shims = os.path.join(ROOT, "baseline", "shims")
    extra = [ref, shims]
    try:
        import matplotlib  # noqa: F401
        extra = [ref]
    except Exception:
        pass
    for path in reversed(extra):
        sys.path.insert(0, path)
    os.environ["PYTHONPATH"] = os.pathsep.join(extra + [os.environ.get("PYTHONPATH", "")])
    # the reference prints progress on stdout; keep stdout for the single JSON line
    sys.stdout.flush()
    saved = os.dup(1)
    _arm_watchdog.fd = saved
    os.dup2(2, 1)
    sampler = ClockSampler(_physical_gpu_index(local_rank), period_s=0.02)
    try:
        import ddlb.cli.benchmark as ddlb_cli
        import tempfile
        tmp = tempfile.mkdtemp(prefix="ddlb_ref_run_")
        sampler.start()
        # headline number with the contract's protocol: ONE event pair around all iterations
        # (the reference's `barrier_at_each_iteration: false` loop, ddlb/benchmark.py:145-159)
        head = _reference_case(ddlb_cli, HEADLINE, args.steps, args.warmup,
                               os.path.join(tmp, "head.csv"), rank, False)
        clocks = sampler.stop()
        # ... and with its default protocol (barrier + synchronize + event pair per iteration)
        head_bar = _reference_case(ddlb_cli, HEADLINE, args.steps, args.warmup,
                                   os.path.join(tmp, "head_bar.csv"), rank, True)
        sec = _reference_case(ddlb_cli, SECONDARY, args.steps, args.warmup,
                              os.path.join(tmp, "sec.csv"), rank, False)
        os.environ["DDLB_MASTER_PORT"] = str(int(os.environ["DDLB_MASTER_PORT"]) + 1)
        e2e = _reference_e2e(HEADLINE, args.steps, args.warmup, rank, world)
    except Exception as exc:
        sys.stdout.flush()
        os.dup2(saved, 1)
        return {"impl": "reference", "unavailable": f"{type(exc).__name__}: {str(exc)[:200]}"}
    sys.stdout.flush()
    os.dup2(saved, 1)
    if rank != 0:
        return {}
    return {
        "metric": "tp_columnwise_tflops", "value": head["tflops"],
        "unit": "TFLOP/s (2*m*n*k/t, whole job)", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "impl": "reference",
        "config": _common_config(args.gpus),
        "implementation": {"name": "reference ddlb.cli.run_benchmark, " + head["label"],
                           "note": "its run() calls torch.cuda.empty_cache() and re-allocates the "
                                   "output every iteration (ddlb/primitives/TPColumnwise/pytorch.py:92)"},
        "protocols": {"back_to_back (value)": head["ms_per_step"],
                      "per_iteration_barrier": head_bar["ms_per_step"]},
        "clocks": clocks, "gpu_launches": 0,
        "e2e": {"value": 2.0 * HEADLINE["m"] * HEADLINE["n"] * HEADLINE["k"] / (e2e["ms_per_step"] * 1e9),
                "unit": "TFLOP/s (2*m*n*k/t, whole job)", "ms_per_step": e2e["ms_per_step"],
                "h2d_bytes_per_step": e2e["h2d_bytes_per_step"],
                "d2h_bytes_per_step": e2e["d2h_bytes_per_step"],
                "mode": "serial: H2D, PyTorchTPColumnwise.run(), D2H on one stream",
                "result_read_back": "this rank's m/N rows of C (the rows of its own input shard)"},
        "secondary": {"metric": "tp_rowwise_tflops",
                      "config": "tp_rowwise GEMM->ReduceScatter m=8192 n=1024 k=8192 fp16",
                      "value": sec["tflops"], "ms_per_step": sec["ms_per_step"]},
    }


def main() -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        result = run_reference(args)
        if result:
            print(json.dumps(result), flush=True)
        return
    # libraries (NCCL's version banner, ...) write to stdout; keep it for the ONE JSON line
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    try:
        result = run_ours(args)
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
    if result:
        print(json.dumps(result), flush=True)


if __name__ == "__main__":
    main()
