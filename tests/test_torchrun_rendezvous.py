"""Regression test for the round-1 reference-arm hang: under torchrun, children that rendezvous
with ``tcp://host:port`` (as the reference's implementations do, ``ddlb/primitives/TPColumnwise/
pytorch.py:53-59``) all become clients of a store nobody hosts unless the ``TORCHELASTIC_*``
variables are removed from their environment. ``bench.py --impl reference`` removes them; this
test runs the same situation on CPU/gloo."""
import os
import subprocess
import sys
import textwrap


def test_tcp_rendezvous_from_torchrun_children_needs_stripped_env(tmp_path):
    script = tmp_path / "child_rdzv.py"
    script.write_text(textwrap.dedent('''
        import datetime, multiprocessing as mp, os

        def child(rank, world, port, q):
            import torch.distributed as dist
            try:
                dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank,
                                        world_size=world, timeout=datetime.timedelta(seconds=60))
                dist.barrier()
                q.put("ok")
                dist.destroy_process_group()
            except Exception as exc:
                q.put("FAIL " + str(exc).splitlines()[0][:120])

        if __name__ == "__main__":
            rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
            assert os.environ.get("TORCHELASTIC_USE_AGENT_STORE") == "True"   # the trap
            for key in [k for k in os.environ if k.startswith("TORCHELASTIC_")]:   # bench.py's fix
                os.environ.pop(key, None)
            ctx = mp.get_context("spawn")
            q = ctx.SimpleQueue()
            p = ctx.Process(target=child, args=(rank, world, int(os.environ["MASTER_PORT"]) + 307, q))
            p.start()
            msg = q.get()
            os.write(1, (f"rank {rank}: {msg}" + chr(10)).encode())       # one write: lines of two ranks never mix
            p.join()
    '''))
    import socket
    with socket.socket() as probe:                 # a port that is free now (and +307 most likely too)
        probe.bind(("127.0.0.1", 0))
        port = probe.getsockname()[1]
    if port + 307 > 65000:
        port -= 1000
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)]
    proc = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    lines = sorted(l for l in proc.stdout.splitlines() if l.startswith("rank "))
    assert lines == ["rank 0: ok", "rank 1: ok"], proc.stdout[-1500:] + proc.stderr[-1500:]


def test_bench_reference_arm_strips_torchelastic_env():
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py")).read()
    assert 'startswith("TORCHELASTIC_")' in src and "os.environ.pop(key, None)" in src
