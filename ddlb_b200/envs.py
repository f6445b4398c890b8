"""Launcher-environment parsing (no CUDA, no torch import).

Covers the reference's launcher variables (Open MPI / SLURM / PMI; reference
``ddlb/envs.py:12-82``) and adds the ``torchrun`` / in-tree launcher variables
(``RANK``, ``LOCAL_RANK``, ``WORLD_SIZE``, ``LOCAL_WORLD_SIZE``), which take
precedence because the GPU box has no ``mpirun``.
"""
from __future__ import annotations

import os
from typing import Iterable, Optional

__all__ = [
    "get_env", "get_rank", "get_local_rank", "get_world_size", "get_local_size",
    "get_master_addr", "get_master_port", "get_device_kind", "rendezvous_port",
]


def get_env(key: str, default: Optional[str] = None,
            fallback_keys: Iterable[str] = ()) -> Optional[str]:
    """Return ``os.environ[key]``, else the first set fallback key, else default."""
    for name in (key, *tuple(fallback_keys)):
        value = os.environ.get(name)
        if value is not None and value != "":
            return value
    return default


def get_rank() -> int:
    return int(get_env("RANK", "0", ("OMPI_COMM_WORLD_RANK", "SLURM_PROCID", "PMI_RANK")))


def get_local_rank() -> int:
    return int(get_env("LOCAL_RANK", "0", ("OMPI_COMM_WORLD_LOCAL_RANK", "SLURM_LOCALID")))


def get_world_size() -> int:
    return int(get_env("WORLD_SIZE", "1", ("OMPI_COMM_WORLD_SIZE", "SLURM_NTASKS", "PMI_SIZE")))


def get_local_size() -> int:
    return int(get_env("LOCAL_WORLD_SIZE", None,
                       ("OMPI_COMM_WORLD_LOCAL_SIZE", "SLURM_NTASKS_PER_NODE"))
               or get_world_size())


def get_master_addr() -> str:
    # 127.0.0.1 rather than "localhost": container hostnames may not resolve.
    return get_env("DDLB_MASTER_ADDR", "127.0.0.1", ("MASTER_ADDR",))


def get_master_port() -> int:
    """Base rendezvous port for benchmark children.

    ``DDLB_MASTER_PORT`` wins. Under torchrun ``MASTER_PORT`` is owned by the
    agent's store, so children rendezvous on an offset from it instead.
    """
    explicit = get_env("DDLB_MASTER_PORT")
    if explicit is not None:
        return int(explicit)
    torchrun_port = get_env("MASTER_PORT")
    if torchrun_port is not None:
        return int(torchrun_port) + 101
    return 12345


def rendezvous_port(sequence: int = 0) -> int:
    """Port for the ``sequence``-th child of a run (identical on every rank).

    The reference reuses one fixed port for every successive child
    (``ddlb/envs.py:75-77``); rotating it avoids lingering-socket collisions.
    """
    return get_master_port() + (int(sequence) % 199)


def get_device_kind() -> str:
    """``"cuda"`` or ``"cpu"`` as requested by ``DDLB_DEVICE`` (default: auto)."""
    return (get_env("DDLB_DEVICE", "auto") or "auto").lower()
