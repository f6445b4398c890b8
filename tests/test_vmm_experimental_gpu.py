"""VMM heap + NVLS multicast window (two real GPUs). Never run on hardware in round 1: opt-in
(``DDLB_RUN_EXPERIMENTAL=1``)."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu,
              pytest.mark.skipif(os.environ.get("DDLB_RUN_EXPERIMENTAL", "0") != "1",
                                 reason="experimental: set DDLB_RUN_EXPERIMENTAL=1")]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_vmm_heap_peer_mapping_and_multicast(tmp_path):
    script = tmp_path / "vmm.py"
    script.write_text(textwrap.dedent(f'''
        import sys
        sys.path.insert(0, {ROOT!r})
        import torch
        from ddlb_b200.communicator import Communicator
        from ddlb_b200.parallel.vmm_heap import VmmHeap
        comm = Communicator()
        comm.init_process_group()
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
