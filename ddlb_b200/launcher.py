"""In-tree multi-process launcher (the GPU box has neither ``mpirun`` nor SLURM).

    python -m ddlb_b200.launcher -np 8 scripts/run_benchmark.py scripts/config.json
    python -m ddlb_b200.launcher -np 2 --cpu -m ddlb_b200.cli.benchmark --primitive ...

Starts ``np`` copies of the command with ``RANK``, ``LOCAL_RANK``, ``WORLD_SIZE``,
``LOCAL_WORLD_SIZE`` and ``DDLB_MASTER_ADDR/PORT`` set (the role ``mpirun -np N``
plays for the reference, ``README.md:80,143``), forwards rank 0's output, and
kills the remaining ranks if one fails. ``torchrun`` works too: ``envs`` reads
its variables.
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time
from typing import List, Optional


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def launch(np_: int, command: List[str], cpu: bool = False, timeout: Optional[float] = None,
           quiet_ranks: bool = True) -> int:
    import tempfile
    port = int(os.environ.get("DDLB_MASTER_PORT", 0)) or _free_port()
    rdzv_dir = tempfile.mkdtemp(prefix="ddlb_rdzv_")
    procs = []
    for rank in range(np_):
        env = dict(os.environ)
        env.update({
            "RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(np_),
            "LOCAL_WORLD_SIZE": str(np_), "DDLB_MASTER_ADDR": "127.0.0.1",
            "DDLB_MASTER_PORT": str(port),
            "DDLB_RDZV_PREFIX": os.path.join(rdzv_dir, "store"),
        })
        env.pop("MASTER_PORT", None)
        if cpu:
            env["DDLB_DEVICE"] = "cpu"
        out = subprocess.DEVNULL if (quiet_ranks and rank != 0) else None
        procs.append(subprocess.Popen(command, env=env, stdout=out,
                                      start_new_session=True))
    deadline = time.monotonic() + timeout if timeout else None
    code = 0
    try:
        pending = set(range(np_))
        while pending:
            for r in sorted(pending):
                rc = procs[r].poll()
                if rc is not None:
                    pending.discard(r)
                    if rc != 0 and code == 0:
                        code = rc
                        print(f"[launcher] rank {r} exited with code {rc}; stopping the others",
                              file=sys.stderr)
                        raise ChildProcessError
            if deadline and time.monotonic() > deadline:
                code = 124
                print("[launcher] timeout", file=sys.stderr)
                raise ChildProcessError
            time.sleep(0.1)
    except (ChildProcessError, KeyboardInterrupt):
        code = code or 130
        for p in procs:                    # exact process groups we started
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except ProcessLookupError:
                    pass
        for p in procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except ProcessLookupError:
                    pass
    import shutil
    shutil.rmtree(rdzv_dir, ignore_errors=True)
    return code


def main(argv: Optional[List[str]] = None) -> None:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawTextHelpFormatter)
    ap.add_argument("-np", "--np", type=int, required=True, help="number of ranks")
    ap.add_argument("--cpu", action="store_true", help="CPU mode (gloo), sets DDLB_DEVICE=cpu")
    ap.add_argument("--timeout", type=float, default=None)
    ap.add_argument("--all-output", action="store_true", help="show every rank's stdout")
    ap.add_argument("-m", dest="module", default=None, help="run a module (python -m)")
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    args = ap.parse_args(argv)
    rest = [a for a in args.rest if a != "--"] if args.rest[:1] == ["--"] else args.rest
    if args.module:
        command = [sys.executable, "-m", args.module, *rest]
    elif rest:
        command = [sys.executable, *rest] if rest[0].endswith(".py") else rest
    else:
        ap.error("nothing to launch")
    sys.exit(launch(args.np, command, cpu=args.cpu, timeout=args.timeout,
                    quiet_ranks=not args.all_output))


if __name__ == "__main__":
    main()
