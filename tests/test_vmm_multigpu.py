"""VMM heap + NVLS multicast window on two real GPUs: peer mapping through POSIX file
descriptors, ``multimem.st`` / ``multimem.ld_reduce`` round trip, and the fused kernels'
``multimem`` transports against the fp32 reference."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _need_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 CUDA devices")


def test_vmm_heap_peer_mapping_and_multicast(tmp_path):
    _need_two_gpus()
    script = tmp_path / "vmm.py"
    script.write_text(textwrap.dedent(f'''
        import sys
        sys.path.insert(0, {ROOT!r})
        import torch
        from ddlb_b200.communicator import Communicator
        from ddlb_b200.parallel.vmm_heap import VmmHeap
        comm = Communicator()
        heap = VmmHeap(comm, 64 << 20)
        off, t = heap.empty((1024,), torch.float32)
        t.fill_(float(comm.rank + 1))
        torch.cuda.synchronize(); heap.barrier(); torch.cuda.synchronize()
        peer = (comm.rank + 1) % comm.world_size
        from ddlb_b200.parallel.symm_heap import _RawCudaBuffer
        raw = _RawCudaBuffer(heap.ptr(off, peer), 4096, heap)
        seen = torch.as_tensor(raw, device=t.device).view(torch.float32)
        assert float(seen[0]) == peer + 1, float(seen[0])          # peer memory is mapped
        ok = heap.multicast_selftest() if heap.has_multicast else None
        print(f"rank {{comm.rank}}: peers ok, multicast={{heap.has_multicast}} selftest={{ok}}", flush=True)
        assert ok in (True, None)
        heap.close(); comm.shutdown()
    '''))
    proc = subprocess.run([sys.executable, "-m", "ddlb_b200.launcher", "-np", "2", "--timeout", "120",
                           "--all-output", str(script)], cwd=ROOT, capture_output=True, text=True,
                          timeout=180)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    assert proc.stdout.count("peers ok") == 2


def test_fused_multimem_transports_two_ranks(tmp_path):
    """Both primitives through ``run_benchmark`` with the NVLS data paths: ``multimem.st`` all-gather
    (both orders) and ``multimem.ld_reduce`` reduce-scatter, validated on a poisoned launch."""
    _need_two_gpus()
    import csv
    import json
    env = dict(os.environ, DDLB_SYMM_MODE="vmm")
    for prim, impls in (
            ("tp_columnwise", {"fused": [{"order": ["AG_before", "AG_after"], "transport": "multimem"}],
                               "fuser": [{"backend": "cuda", "multicast_protocol": "multimem"}]}),
            ("tp_rowwise", {"fused": [{"transport": "multimem"}],
                            "fuser": [{"backend": "cuda", "multicast_protocol": "multimem"}]})):
        out_csv = tmp_path / f"{prim}.csv"
        cfg = {"benchmark": {"primitive": prim, "m": 8192, "n": 1024, "k": 2048, "dtype": "bfloat16",
                             "validate": True, "num_iterations": 10, "num_warmups": 3,
                             "time_measurement_backend": "cuda_event", "output_csv": str(out_csv),
                             "implementations": impls}}
        cfg_path = tmp_path / f"{prim}.json"
        cfg_path.write_text(json.dumps(cfg))
        proc = subprocess.run([sys.executable, "-m", "ddlb_b200.launcher", "-np", "2", "--timeout", "300",
                               os.path.join(ROOT, "scripts", "run_benchmark.py"), str(cfg_path)],
                              cwd=ROOT, capture_output=True, text=True, timeout=400, env=env)
        assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
        rows = list(csv.DictReader(open(out_csv)))
        assert len(rows) == (3 if prim == "tp_columnwise" else 2)
        for r in rows:
            assert r["valid"] == "True" and r["error"] == "", r
        assert any(r["implementation"].startswith("fused[as fuser]") and
                   "multicast_protocol=multimem" in r["implementation"] for r in rows)
