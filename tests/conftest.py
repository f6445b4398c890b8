import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs at least two CUDA devices")


def _cuda_count():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    ndev = None
    for item in items:
        if "gpu" in item.keywords or "multigpu" in item.keywords:
            if ndev is None:
                ndev = _cuda_count()
            if ndev == 0:
                item.add_marker(pytest.mark.skip(reason="no CUDA device"))
            elif "multigpu" in item.keywords and ndev < 2:
                item.add_marker(pytest.mark.skip(reason="needs >= 2 CUDA devices"))


@pytest.fixture
def clean_env(monkeypatch):
    for key in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "OMPI_COMM_WORLD_RANK",
                "OMPI_COMM_WORLD_LOCAL_RANK", "OMPI_COMM_WORLD_SIZE", "OMPI_COMM_WORLD_LOCAL_SIZE",
                "SLURM_PROCID", "SLURM_LOCALID", "SLURM_NTASKS", "SLURM_NTASKS_PER_NODE",
                "PMI_RANK", "PMI_SIZE", "DDLB_MASTER_ADDR", "DDLB_MASTER_PORT", "MASTER_ADDR",
                "MASTER_PORT", "DDLB_DEVICE"):
        monkeypatch.delenv(key, raising=False)
    return monkeypatch
