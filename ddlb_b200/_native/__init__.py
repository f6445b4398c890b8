"""ctypes binding of ``libddlb_b200.so`` (C ABI in ``include/ddlb_b200.h``).

The library is built in-tree (``python -m ddlb_b200._native.build``) and loaded on
first use. If it is missing the load fails loudly: there is no silent PyTorch
fallback for the product kernels.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import (POINTER, Structure, byref, c_char_p, c_int, c_int32, c_int64, c_uint32,
                    c_void_p)
from typing import Optional

MAX_WORLD = 8
TRACE_ROLES, TRACE_SLOTS = 4, 128
IPC_HANDLE_BYTES = 64
MODE_PLAIN, MODE_AG_BEFORE, MODE_AG_AFTER, MODE_RS = 0, 1, 2, 3
DTYPE_F16, DTYPE_BF16, DTYPE_F32 = 0, 1, 2
B_KN, B_NK = 0, 1
AG_PULL, AG_PUSH, AG_PUSH_MULTICAST = 0, 1, 2
EPI_TMA, EPI_ST128, EPI_MULTICAST = 0, 1, 2
RS_PUSH, RS_PULL, RS_PULL_MULTICAST = 0, 1, 2
FP32_3XTF32, FP32_TF32 = 0, 1
ACT_NONE, ACT_RELU, ACT_SILU, ACT_GELU_TANH = 0, 1, 2, 3

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libddlb_b200.so")
_lib: Optional[ctypes.CDLL] = None


class NativeError(RuntimeError):
    pass


class TpGemmConfig(Structure):
    """Mirror of ``ddlb_tp_gemm_config`` (``include/ddlb_b200.h``)."""
    _fields_ = [
        ("mode", c_int32), ("dtype", c_int32), ("b_layout", c_int32), ("cta_group", c_int32),
        ("block_n", c_int32), ("group_m", c_int32), ("num_sms", c_int32),
        ("rank", c_int32), ("world", c_int32), ("chunk_rows", c_int32),
        ("ag_transport", c_int32), ("epi_transport", c_int32), ("rs_transport", c_int32),
        ("unrotated", c_int32), ("fp32_mode", c_int32), ("epilogue_act", c_int32),
        ("comm_warps", c_int32), ("cluster_pairs", c_int32), ("pdl", c_int32), ("wide_n", c_int32),
        ("m", c_int64), ("n", c_int64), ("k", c_int64),
        ("a", c_void_p), ("b", c_void_p),
        ("out", c_void_p * MAX_WORLD), ("peer_a", c_void_p * MAX_WORLD),
        ("stage", c_void_p * MAX_WORLD), ("mc_ptr", c_void_p),
        ("sig_local", c_void_p), ("sig_peer", c_void_p * MAX_WORLD),
        ("trace", c_void_p),
    ]


def lib() -> ctypes.CDLL:
    """Load (building first if the sources are newer and nvcc is available)."""
    global _lib
    if _lib is not None:
        return _lib
    if os.environ.get("DDLB_NO_AUTOBUILD", "0") != "1":
        try:
            from . import build as _build
            if _build.is_stale():
                _build.build()
        except Exception as exc:  # no nvcc on the box, or a compile error
            if not os.path.exists(LIB_PATH):
                raise NativeError(f"libddlb_b200.so is missing and could not be built: {exc}")
            # an OLD library would run old kernels (or miss new symbols) without a word
            if os.environ.get("DDLB_ALLOW_STALE_LIB", "0") != "1":
                raise NativeError(
                    "csrc/ or include/ changed since libddlb_b200.so was built and the rebuild "
                    f"failed ({str(exc).splitlines()[0][:200]}); fix the build or set "
                    "DDLB_ALLOW_STALE_LIB=1 to load the stale library") from exc
            import warnings
            warnings.warn(f"loading a STALE libddlb_b200.so (rebuild failed: {exc})")
    if not os.path.exists(LIB_PATH):
        raise NativeError(
            f"{LIB_PATH} not found; run `python -m ddlb_b200._native.build` (needs nvcc)")
    L = ctypes.CDLL(LIB_PATH)
    L.ddlb_last_error.restype = c_char_p
    L.ddlb_version.restype = c_int
    L.ddlb_topology_json.argtypes = [c_char_p, c_int64, POINTER(c_int64)]
    L.ddlb_symm_create.argtypes = [c_int, c_int, c_int, c_int64, POINTER(c_void_p)]
    L.ddlb_symm_export.argtypes = [c_void_p, c_void_p]
    L.ddlb_symm_open_peers.argtypes = [c_void_p, c_void_p]
    L.ddlb_symm_adopt_peer.argtypes = [c_void_p, c_int, c_void_p]
    L.ddlb_symm_base.argtypes = [c_void_p, c_int]
    L.ddlb_symm_base.restype = c_void_p
    L.ddlb_symm_bytes.argtypes = [c_void_p]
    L.ddlb_symm_bytes.restype = c_int64
    L.ddlb_symm_barrier.argtypes = [c_void_p, c_void_p]
    L.ddlb_symm_destroy.argtypes = [c_void_p]
    L.ddlb_tp_gemm_sig_words.argtypes = [POINTER(TpGemmConfig)]
    L.ddlb_tp_gemm_sig_words.restype = c_int64
    L.ddlb_tp_gemm_create.argtypes = [POINTER(TpGemmConfig), POINTER(c_void_p)]
    L.ddlb_tp_gemm_run.argtypes = [c_void_p, c_uint32, c_void_p]
    L.ddlb_tp_gemm_info.argtypes = [c_void_p, POINTER(c_int32 * 12)]
    L.ddlb_tp_gemm_destroy.argtypes = [c_void_p]
    L.ddlb_tp_gemm_run_emulated.argtypes = [POINTER(c_void_p), c_int, c_uint32, c_void_p]
    L.ddlb_tp_gemm_run_emulated_static.argtypes = [POINTER(c_void_p), c_int, c_uint32, c_void_p]
    L.ddlb_debug_schedule.argtypes = [POINTER(TpGemmConfig), POINTER(c_int32 * 12)]
    L.ddlb_debug_tile_coords.argtypes = [POINTER(TpGemmConfig), c_int32, c_int32, POINTER(c_int32)]
    L.ddlb_debug_pull_units.argtypes = [POINTER(TpGemmConfig), c_int64, POINTER(c_int64), POINTER(c_int64)]
    c_int_p = POINTER(c_int)
    L.ddlb_vmm_multicast_supported.argtypes = [c_int, c_int_p]
    L.ddlb_vmm_create.argtypes = [c_int, c_int, c_int, c_int64, POINTER(c_void_p)]
    L.ddlb_vmm_export_fd.argtypes = [c_void_p, c_int_p]
    L.ddlb_vmm_import_peer.argtypes = [c_void_p, c_int, c_int]
    L.ddlb_vmm_mc_create.argtypes = [c_void_p, c_int_p]
    L.ddlb_vmm_mc_import.argtypes = [c_void_p, c_int]
    L.ddlb_vmm_mc_add_device.argtypes = [c_void_p]
    L.ddlb_vmm_mc_bind.argtypes = [c_void_p]
    L.ddlb_vmm_base.argtypes = [c_void_p, c_int]
    L.ddlb_vmm_base.restype = c_void_p
    L.ddlb_vmm_mc_base.argtypes = [c_void_p]
    L.ddlb_vmm_mc_base.restype = c_void_p
    L.ddlb_vmm_bytes.argtypes = [c_void_p]
    L.ddlb_vmm_bytes.restype = c_int64
    L.ddlb_vmm_selftest_store.argtypes = [c_void_p, c_int64, c_int, c_void_p]
    L.ddlb_vmm_selftest_reduce.argtypes = [c_void_p, c_int64, c_void_p, c_int, c_void_p]
    L.ddlb_vmm_destroy.argtypes = [c_void_p]
    L.ddlb_vmm_has_multicast.argtypes = [c_void_p]
    L.ddlb_barrier_raw.argtypes = [POINTER(c_void_p), c_int, c_int, c_uint32, c_void_p]
    L.ddlb_copy.argtypes = [c_void_p, c_void_p, c_int64, c_int, c_void_p]
    L.ddlb_copy_pattern.argtypes = [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]
    L.ddlb_max_active_clusters.argtypes = [c_int, c_int, POINTER(c_int)]
    L.ddlb_read_error_word.argtypes = [POINTER(c_uint32 * 4)]
    _lib = L
    return L


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().ddlb_last_error().decode(errors="replace")
        raise NativeError(f"{what}: {msg}" if what else msg)


def loaded() -> bool:
    return _lib is not None


def topology_json() -> str:
    L = lib()
    need = c_int64(0)
    check(L.ddlb_topology_json(None, 0, byref(need)), "topology")
    buf = ctypes.create_string_buffer(need.value + 16)
    check(L.ddlb_topology_json(buf, len(buf), byref(need)), "topology")
    return buf.value.decode()


def read_error_word():
    word = (c_uint32 * 4)()
    check(lib().ddlb_read_error_word(byref(word)), "read_error_word")
    return list(word)
