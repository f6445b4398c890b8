"""Benchmark core: one spawned child per implementation, timed on the device.

Behavioural contract follows the reference's ``ddlb/benchmark.py``:

* the parent never touches CUDA; rank comes from the launcher environment
  (``benchmark.py:17,331``);
* every implementation runs in a freshly *spawned* child so CUDA context,
  process group and library state never leak between implementations
  (``benchmark.py:336-370``);
* protocol inside the child (``benchmark.py:82-204``): warm-ups, a
  ``cudaProfilerStart/Stop`` bracket around 5 runs, warm-ups again, the timed
  loop in one of four modes, element-wise MAX over ranks, mean/std/min/max and
  ``2mnk/t`` TFLOP/s;
* rank 0 appends each row to the CSV as soon as it arrives
  (``benchmark.py:372-384``), same column order, extra columns after it.

Deliberate differences: the child always posts a row (``valid=False`` plus an
``error`` column) and the parent polls ``is_alive`` so a crashed child cannot
hang the run; cross-rank sync before each timed iteration uses the
device barrier on the symmetric signal pads when the implementation has one,
not a dummy NCCL all-reduce; ``size`` is kept in the ``option`` column (Q3).
"""
from __future__ import annotations

import csv
import multiprocessing as mp
import os
import queue as _queue
import socket
import time
import traceback
from typing import Any, Dict, List, Optional

from .envs import get_rank, get_world_size
from .primitives.registry import ALLOWED_PRIMITIVES, base_impl_name, load_impl_class

TIME_BACKENDS = ("cpu_clock", "cuda_event")
PROFILED_RUNS = 5

# Column order of the reference's row (ddlb/benchmark.py:220-243), then ours.
REFERENCE_COLUMNS = [
    "implementation", "mean_time (ms)", "std_time", "min_time", "max_time",
    "m", "n", "k", "dtype", "Throughput (TFLOPS)", "Throughput std (TFLOPS)",
    "world_size", "hostname", "time_measurement_backend",
    "barrier_at_each_iteration", "option",
]
EXTRA_COLUMNS = ["roofline_ms", "pct_roofline", "flops_per_gpu", "comm_bytes_per_gpu", "error"]

_child_sequence = 0


def _merged_options(impl_class, impl_opts: Dict[str, Any]) -> Dict[str, Any]:
    """Defaults overlaid with the given options; unknown keys are dropped
    silently, like the reference's worker (``benchmark.py:76-77``)."""
    options = dict(getattr(impl_class, "DEFAULT_OPTIONS", {}))
    options.update({k: v for k, v in impl_opts.items() if k in options})
    return options


def describe_implementation(impl_class, base_impl: str, options: Dict[str, Any]):
    """``(label, option_string)`` for a result row.

    The label excludes ``size`` exactly like the reference
    (``benchmark.py:106-119``); the option string keeps it. Classes that serve a
    reference implementation NAME with the native kernel (``SERVED_BY``, see
    ``primitives/TPColumnwise/compat.py``) are labelled ``fused[as <name>]`` and list the
    options that have no equivalent under ``ignored:`` so a results file never attributes
    native-kernel timings to nvFuser / Transformer Engine / JAX.
    """
    keys = [k for k in getattr(impl_class, "DEFAULT_OPTIONS", {}) if k in options]
    ignored = []
    honoured = getattr(impl_class, "honoured_options", None)
    if honoured is not None:
        keep = honoured(options)
        ignored = [k for k in keys if k not in keep]
        keys = [k for k in keys if k in keep]
    shown = [k for k in keys if k != "size"]
    served_by = getattr(impl_class, "SERVED_BY", None)
    label = f"{served_by}[as {base_impl}]" if served_by else base_impl
    if shown:
        label += " (" + ", ".join(f"{k}={options[k]}" for k in shown) + ")"
    option_str = ", ".join(f"{k}={options[k]}" for k in keys)
    if ignored:
        option_str += ("; " if option_str else "") + "ignored: " + ",".join(
            f"{k}={options[k]}" for k in ignored)
    return label, option_str


def _empty_row(label, option_str, m, n, k, dtype, backend, barrier, validate, error):
    """A result row with the reference's columns in the reference's order
    (``REFERENCE_COLUMNS``), ``valid`` when validating, then ``EXTRA_COLUMNS``."""
    row = {
        "implementation": label, "mean_time (ms)": float("nan"), "std_time": float("nan"),
        "min_time": float("nan"), "max_time": float("nan"), "m": m, "n": n, "k": k,
        "dtype": dtype, "Throughput (TFLOPS)": 0.0, "Throughput std (TFLOPS)": 0.0,
        "world_size": get_world_size(), "hostname": socket.gethostname(),
        "time_measurement_backend": backend, "barrier_at_each_iteration": barrier,
        "option": option_str,
    }
    if validate:
        row["valid"] = False
    row.update({"roofline_ms": float("nan"), "pct_roofline": float("nan"),
                "flops_per_gpu": float("nan"), "comm_bytes_per_gpu": float("nan"),
                "error": error})
    expected = REFERENCE_COLUMNS + (["valid"] if validate else []) + EXTRA_COLUMNS
    assert list(row) == expected, "result-row schema drifted from the documented column order"
    return row


class _Range:
    """NVTX range around a benchmark phase (no-op on CPU); shows up in Nsight timelines
    next to the ``cudaProfilerStart/Stop`` bracket. The reference emits none."""

    def __init__(self, comm, name: str):
        self._on = bool(comm.is_cuda)
        self._name = name

    def __enter__(self):
        if self._on:
            try:
                import torch
                torch.cuda.nvtx.range_push(self._name)
            except Exception:
                self._on = False
        return self

    def __exit__(self, *exc):
        if self._on:
            import torch
            torch.cuda.nvtx.range_pop()


class _L2Flusher:
    """Writes a buffer larger than L2 between timed iterations (``DDLB_L2_FLUSH=1``), for
    shapes whose working set would otherwise stay L2-resident from one run to the next."""

    def __init__(self, comm):
        self.buf = None
        if comm.is_cuda and os.environ.get("DDLB_L2_FLUSH", "0") == "1":
            import torch
            self.buf = torch.empty(256 << 20, dtype=torch.uint8, device=comm.device)

    def __call__(self) -> None:
        if self.buf is not None:
            self.buf.zero_()


def _sync(comm) -> None:
    if comm.is_cuda:
        import torch
        torch.cuda.synchronize()


def _device_barrier(impl) -> None:
    """Cross-rank sync before a timed iteration.

    With a symmetric heap the barrier is a kernel on the signal pads, enqueued on the
    timing stream: no host synchronisation, so neither launch latency nor host-side skew
    between the rank processes is charged to the iteration that follows. Without one
    (CPU mode) it is the reference's dummy all-reduce + synchronize
    (``ddlb/benchmark.py:134-137``).
    """
    comm = impl.communicator
    if comm.world_size == 1:
        return
    if comm.is_cuda:
        comm.symmetric_heap().barrier()
        return
    if comm.pg_initialized:
        import torch
        import torch.distributed as dist
        dummy = torch.zeros(1, device=comm.device, dtype=torch.int32)
        dist.all_reduce(dummy)
        _sync(comm)


def time_implementation(impl, num_iterations: int, backend: str, barrier: bool):
    """Run the timed loop; returns ``(per-iteration ms list, last result)``."""
    comm = impl.communicator
    if backend not in TIME_BACKENDS:
        raise ValueError(f"Unknown time_measurement_backend: {backend}")
    if backend == "cuda_event" and not comm.is_cuda:
        backend = "cpu_clock"           # CPU mode has no events
    last = None
    times: List[float] = []
    if num_iterations <= 0:
        return times, last

    if backend == "cuda_event":
        import torch
        if barrier:
            starts = [torch.cuda.Event(enable_timing=True) for _ in range(num_iterations)]
            ends = [torch.cuda.Event(enable_timing=True) for _ in range(num_iterations)]
            if comm.world_size > 1:
                comm.symmetric_heap()          # every implementation gets the device barrier
            comm.barrier()
            flush = _L2Flusher(comm)
            for i in range(num_iterations):
                flush()
                _device_barrier(impl)
                starts[i].record()
                last = impl.run()
                ends[i].record()
            torch.cuda.synchronize()
            times = [starts[i].elapsed_time(ends[i]) for i in range(num_iterations)]
        else:
            start = torch.cuda.Event(enable_timing=True)
            end = torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            start.record()
            for _ in range(num_iterations):
                last = impl.run()
            end.record()
            torch.cuda.synchronize()
            times = [start.elapsed_time(end) / num_iterations] * num_iterations
    else:
        if barrier:
            for _ in range(num_iterations):
                comm.barrier()
                t0 = time.perf_counter()
                last = impl.run()
                _sync(comm)
                times.append((time.perf_counter() - t0) * 1e3)
        else:
            _sync(comm)
            t0 = time.perf_counter()
            for _ in range(num_iterations):
                last = impl.run()
            _sync(comm)
            times = [(time.perf_counter() - t0) * 1e3 / num_iterations] * num_iterations
    return times, last


def _benchmark_worker_entry(impl_id: str, m: int, n: int, k: int, dtype: str,
                            num_warmups: int, num_iterations: int, impl_opts: Dict[str, Any],
                            validate: bool, result_queue,
                            time_measurement_backend: str = "cpu_clock",
                            barrier_at_each_iteration: bool = True,
                            primitive: str = "tp_rowwise") -> None:
    """Child entry point: build, warm up, profile-bracket, time, validate, post a row."""
    import numpy as np

    base_impl = base_impl_name(impl_id)
    label, option_str = base_impl, ""
    impl = None
    row: Optional[Dict[str, Any]] = None
    try:
        impl_class = load_impl_class(primitive, base_impl)
        options = _merged_options(impl_class, impl_opts)
        label, option_str = describe_implementation(impl_class, base_impl, options)

        impl = impl_class(m=m, n=n, k=k, dtype=dtype, **options)
        comm = impl.communicator

        with _Range(comm, f"ddlb/{impl_id}/warmup"):
            for _ in range(num_warmups):
                impl.run()
            _sync(comm)

        cudart = None
        if comm.is_cuda:
            import torch
            try:
                cudart = torch.cuda.cudart()
                cudart.cudaProfilerStart()
            except Exception:
                cudart = None
        with _Range(comm, f"ddlb/{impl_id}/profiled"):
            for _ in range(PROFILED_RUNS):
                impl.run()
            _sync(comm)
        if cudart is not None:
            try:
                cudart.cudaProfilerStop()
            except Exception:
                pass

        for _ in range(num_warmups):
            impl.run()

        with _Range(comm, f"ddlb/{impl_id}/timed"):
            times, last_result = time_implementation(
                impl, num_iterations, time_measurement_backend, barrier_at_each_iteration)
        times = comm.all_reduce_max(times) if times else times

        row = _empty_row(label, option_str, m, n, k, dtype, time_measurement_backend,
                         barrier_at_each_iteration, validate, "")
        if times:
            thr_const = (2.0 * m * n * k) / 1e9        # ms -> s, FLOP -> TFLOP
            thr = [thr_const / t for t in times if t > 0]
            row.update({
                "mean_time (ms)": float(np.mean(times)), "std_time": float(np.std(times)),
                "min_time": float(min(times)), "max_time": float(max(times)),
                "Throughput (TFLOPS)": float(np.mean(thr)) if thr else 0.0,
                "Throughput std (TFLOPS)": float(np.std(thr)) if thr else 0.0,
            })
            from .utils.roofline import roofline_ms
            flops, comm_bytes = impl.flops_per_gpu(), impl.comm_bytes_per_gpu()
            roof = roofline_ms(flops, comm_bytes, dtype)
            row.update({"roofline_ms": roof, "flops_per_gpu": flops,
                        "comm_bytes_per_gpu": comm_bytes,
                        "pct_roofline": 100.0 * roof / row["mean_time (ms)"]})
        if validate:
            # never validate a launch that ran on the same bytes as the launches before it: a
            # tile consumed before it arrived would read identical stale data and pass. Poison
            # what must cross the wire, run once more, check that run (strictly for the fused
            # kernels; the library baselines keep the reference's tolerance).
            _sync(comm)
            impl.poison()
            last_result = impl.run()
            _sync(comm)
            try:
                impl.validate(last_result, strict=hasattr(impl, "op"))
                row["valid"] = True
            except Exception as exc:       # mismatch is a result, not a crash
                row["valid"] = False
                row["error"] = f"validation: {str(exc).splitlines()[0][:200]}"
                print(f"Warning: Validation failed for {impl_id} with error: {exc}")
    except BaseException as exc:           # always post a row
        traceback.print_exc()
        row = _empty_row(label, option_str, m, n, k, dtype, time_measurement_backend,
                         barrier_at_each_iteration, validate,
                         f"{type(exc).__name__}: {str(exc).splitlines()[0][:300] if str(exc) else ''}")
    finally:
        result_queue.put(row)
        try:
            if impl is not None:
                impl.close()
        except Exception:
            pass
        try:
            from .communicator import Communicator
            Communicator().shutdown()
        except Exception:
            pass


def append_row_csv(path: str, row: Dict[str, Any]) -> None:
    """Append one row; header only when the file is empty (reference :375-384)."""
    directory = os.path.dirname(path)
    if directory:
        os.makedirs(directory, exist_ok=True)
    try:
        write_header = os.path.getsize(path) == 0
    except OSError:
        write_header = True
    with open(path, "a", newline="") as fh:
        writer = csv.DictWriter(fh, fieldnames=list(row.keys()), quoting=csv.QUOTE_MINIMAL)
        if write_header:
            writer.writeheader()
        writer.writerow(row)


def _rows_already_done(path: Optional[str]) -> set:
    done = set()
    if not path or not os.path.exists(path):
        return done
    with open(path, newline="") as fh:
        for r in csv.DictReader(fh):
            if (r.get("error") or "") == "":
                done.add((r.get("implementation"), r.get("option"), r.get("m"), r.get("n"),
                          r.get("k"), r.get("dtype"), r.get("world_size")))
    return done


class PrimitiveBenchmarkRunner:
    """Run a list of implementations of one primitive at one shape.

    Constructor arguments match the reference's runner
    (``ddlb/benchmark.py:269-317``); ``resume`` and ``child_timeout`` are new.
    """

    ALLOWED_PRIMITIVES = set(ALLOWED_PRIMITIVES)

    def __init__(self, primitive: str, m: int, n: int, k: int, implementations: List[str],
                 dtype: str = "float32", validate: bool = True, num_iterations: int = 5,
                 num_warmups: int = 2,
                 implementation_options: Optional[Dict[str, Dict]] = None,
                 output_csv: Optional[str] = None,
                 time_measurement_backend: str = "cpu_clock",
                 barrier_at_each_iteration: bool = True,
                 resume: bool = False, child_timeout: Optional[float] = None):
        if primitive not in self.ALLOWED_PRIMITIVES:
            raise ValueError(f"Unknown primitive: {primitive}")
        if time_measurement_backend not in TIME_BACKENDS:
            raise ValueError(f"Unknown time_measurement_backend: {time_measurement_backend}")
        self.primitive = primitive
        self.m, self.n, self.k = m, n, k
        self.implementations = list(implementations)
        self.dtype = dtype
        self.validate = validate
        self.num_iterations = num_iterations
        self.num_warmups = num_warmups
        self.implementation_options = implementation_options or {}
        self.output_csv = output_csv
        self.time_measurement_backend = time_measurement_backend
        self.barrier_at_each_iteration = barrier_at_each_iteration
        self.resume = resume
        self.child_timeout = float(child_timeout if child_timeout is not None
                                   else os.environ.get("DDLB_CHILD_TIMEOUT", 1800))

    def _run_child(self, ctx, impl_id: str, impl_opts: Dict[str, Any]) -> Dict[str, Any]:
        global _child_sequence
        _child_sequence += 1
        os.environ["DDLB_CHILD_SEQUENCE"] = str(_child_sequence)   # inherited by spawn
        result_queue = ctx.Queue()
        proc = ctx.Process(
            target=_benchmark_worker_entry,
            args=(impl_id, self.m, self.n, self.k, self.dtype, self.num_warmups,
                  self.num_iterations, impl_opts, self.validate, result_queue,
                  self.time_measurement_backend, self.barrier_at_each_iteration,
                  self.primitive))
        proc.start()
        deadline = time.monotonic() + self.child_timeout
        row, error = None, None
        while row is None:
            try:
                row = result_queue.get(timeout=0.5)
            except _queue.Empty:
                if not proc.is_alive():
                    try:                      # a row may have landed just before exit
                        row = result_queue.get(timeout=0.5)
                    except _queue.Empty:
                        error = f"child exited with code {proc.exitcode} before posting a result"
                        break
                elif time.monotonic() > deadline:
                    error = f"child timed out after {self.child_timeout:.0f} s"
                    proc.terminate()
                    break
        proc.join(timeout=30)
        if proc.is_alive():
            proc.kill()
            proc.join()
        if row is None:
            base = base_impl_name(impl_id)
            row = _empty_row(base, "", self.m, self.n, self.k, self.dtype,
                             self.time_measurement_backend, self.barrier_at_each_iteration,
                             self.validate, error or "no result")
        return row

    def run(self):
        import pandas as pd

        rank = get_rank()
        ctx = mp.get_context("spawn")
        csv_path = None
        if rank == 0 and self.output_csv and str(self.output_csv).strip():
            csv_path = self.output_csv
            directory = os.path.dirname(csv_path)
            if directory:
                os.makedirs(directory, exist_ok=True)
            open(csv_path, "a").close()
        # every rank must take the same skip decisions, so resume reads the
        # file on every rank (single node: shared filesystem)
        done = _rows_already_done(self.output_csv) if self.resume else set()

        iterator = self.implementations
        if rank == 0:
            try:
                from tqdm import tqdm
                iterator = tqdm(self.implementations, desc="Running benchmarks", position=0)
            except Exception:
                pass

        results = []
        for impl_id in iterator:
            impl_opts = self.implementation_options.get(impl_id, {})
            if done:
                try:
                    cls = load_impl_class(self.primitive, base_impl_name(impl_id))
                    label, opt = describe_implementation(
                        cls, base_impl_name(impl_id), _merged_options(cls, impl_opts))
                    key = (label, opt, str(self.m), str(self.n), str(self.k), self.dtype,
                           str(get_world_size()))
                    if key in done:
                        if rank == 0:
                            print(f"Skipping {impl_id}: already in {self.output_csv}")
                        continue
                except Exception:
                    pass
            if rank == 0:
                print(f"Running benchmark for {impl_id} with options {impl_opts}")
            row = self._run_child(ctx, impl_id, impl_opts)
            if rank == 0:
                print(pd.DataFrame([row]).to_string(index=False))
                if csv_path:
                    append_row_csv(csv_path, row)
            results.append(row)
        return pd.DataFrame(results)

    def plot_results(self, results=None, path: Optional[str] = None) -> Optional[str]:
        """Bar chart of mean time +- std, written to a file (headless).

        The reference opens an interactive window (``benchmark.py:391-425``);
        there is no display here, so the chart is saved instead: a PNG through
        matplotlib's Agg backend when matplotlib is importable, else an SVG.
        """
        from .utils.plot import bar_chart
        if results is None:
            results = self.run()
        if get_rank() != 0 or len(results) == 0:
            return None
        title = (f"{self.primitive.upper()} Benchmark  Size: ({self.m},{self.n},{self.k}), "
                 f"Dtype: {self.dtype}")
        path = path or f"results/{self.primitive}_{self.m}x{self.k}x{self.n}_{self.dtype}"
        return bar_chart(list(results["implementation"]), list(results["mean_time (ms)"]),
                         list(results["std_time"]), title, "Time (ms)", path)
