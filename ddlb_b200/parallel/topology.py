"""NVLink / NVSwitch topology report.

CUDA-side facts (peer access matrix, P2P attributes, multicast support) come from
the native library (``ddlb_topology_json``); NVLink link states come from NVML
through ``pynvml`` when it is importable. The reference's communicator only
selects a device (``ddlb/communicator.py:56-60``).
"""
from __future__ import annotations

import json
from typing import Any, Dict


def nvlink_summary() -> Dict[str, Any]:
    out: Dict[str, Any] = {"available": False}
    try:
        import pynvml
        pynvml.nvmlInit()
    except Exception as exc:
        out["error"] = str(exc)
        return out
    try:
        devices = []
        for i in range(pynvml.nvmlDeviceGetCount()):
            h = pynvml.nvmlDeviceGetHandleByIndex(i)
            active = 0
            for link in range(18):
                try:
                    if pynvml.nvmlDeviceGetNvLinkState(h, link) == pynvml.NVML_FEATURE_ENABLED:
                        active += 1
                except Exception:
                    break
            devices.append({"index": i, "nvlink_active_links": active})
        out.update({"available": True, "devices": devices})
    finally:
        try:
            pynvml.nvmlShutdown()
        except Exception:
            pass
    return out


def topology_report(comm=None) -> Dict[str, Any]:
    from .. import _native as N
    report = json.loads(N.topology_json())
    report["nvlink"] = nvlink_summary()
    if comm is not None:
        report["rank"] = comm.rank
        report["world_size"] = comm.world_size
        report["local_rank"] = comm.local_rank
    devs = report.get("devices", [])
    report["all_peers_accessible"] = bool(devs) and all(
        all(p["access"] == 1 for p in d["peers"]) for d in devs)
    report["multicast_supported"] = bool(devs) and all(d["multicast_supported"] == 1 for d in devs)
    report["max_active_clusters"] = max_active_clusters()
    return report


def max_active_clusters() -> Dict[str, int]:
    """Resident clusters of the flagship GEMM kernel per cluster size on the current device
    (clusters are confined to a GPC, so 148 SMs need not give 148 / size of them)."""
    import ctypes
    from .. import _native as N
    out = {}
    try:
        lib = N.lib()
        for size in (1, 2, 4, 8):
            n = ctypes.c_int(0)
            if lib.ddlb_max_active_clusters(size, 256, ctypes.byref(n)) == 0:
                out[str(size)] = int(n.value)
    except Exception:
        pass
    return out


if __name__ == "__main__":
    print(json.dumps(topology_report(), indent=1))
