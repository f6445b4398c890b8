"""Run under ``python -m ddlb_b200.launcher -np N``: 10 000 device barriers back to back with
the epoch counter starting just below 2^32 (the flag comparison is wrap-safe), interleaved with
a token each rank writes for its right neighbour, which must always be visible after the
barrier. Used by tests/test_multigpu.py."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from ddlb_b200 import _native as N  # noqa: E402
from ddlb_b200.communicator import Communicator  # noqa: E402
from ddlb_b200.parallel.symm_heap import _RawCudaBuffer  # noqa: E402


def main():
    comm = Communicator()
    d, r = comm.world_size, comm.rank
    heap = comm.symmetric_heap(64 << 20)
    lib = N.lib()
    off, mine = heap.empty((16,), torch.int64)
    mine.zero_()
    torch.cuda.synchronize()
    heap.barrier()
    torch.cuda.synchronize()
    right = (r + 1) % d
    raw = _RawCudaBuffer(heap.ptr(off, right), 16 * 8, heap)
    peer = torch.as_tensor(raw, device=mine.device).view(torch.int64)      # right neighbour's slot
    # separate pads so that the heap's own barrier counter is not disturbed: a second signal area
    sig_off = heap.alloc_signals(64)
    pads = (ctypes.c_void_p * d)(*[heap.ptr(sig_off, p) for p in range(d)])
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    start = (1 << 32) - 5000                               # wraps after 5000 of the 10000 barriers
    # the comparison is wrap-safe for flags within 2^31 of their target: bring the (zeroed)
    # arrival words to "epoch start-1 reached" on every rank before the first barrier
    heap.tensor(sig_off, (64,), torch.int32).fill_(start - 1 - (1 << 32))
    torch.cuda.synchronize()
    heap.barrier()
    torch.cuda.synchronize()
    bad = 0
    for i in range(10000):
        epoch = (start + i) & 0xFFFFFFFF
        peer[0] = i + 1                                    # remote store, then the barrier
        N.check(lib.ddlb_barrier_raw(pads, r, d, epoch, stream), "barrier_raw")
        if i % 500 == 499 or i == 4999 or i == 5000:
            torch.cuda.synchronize()
            if int(mine[0].item()) < i + 1:                # left neighbour's token of this round
                bad += 1
    torch.cuda.synchronize()
    assert bad == 0, f"rank {r}: {bad} barriers let a stale token through"
    comm.barrier()
    print(f"barrier stress ok rank {r}", flush=True)
    comm.shutdown()


if __name__ == "__main__":
    main()
