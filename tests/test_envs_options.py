import os

import pytest

from ddlb_b200 import envs
from ddlb_b200.primitives.options import EnvVarGuard, OptionsManager, setup_ucc_env_vars


def test_defaults(clean_env):
    assert envs.get_rank() == 0 and envs.get_world_size() == 1
    assert envs.get_local_rank() == 0 and envs.get_local_size() == 1
    assert envs.get_master_addr() == "127.0.0.1"
    assert envs.get_master_port() == 12345


def test_fallback_order(clean_env):
    clean_env.setenv("PMI_RANK", "5")
    clean_env.setenv("PMI_SIZE", "9")
    assert envs.get_rank() == 5 and envs.get_world_size() == 9
    clean_env.setenv("SLURM_PROCID", "4")
    clean_env.setenv("SLURM_NTASKS", "8")
    clean_env.setenv("SLURM_LOCALID", "2")
    clean_env.setenv("SLURM_NTASKS_PER_NODE", "4")
    assert envs.get_rank() == 4 and envs.get_world_size() == 8
    assert envs.get_local_rank() == 2 and envs.get_local_size() == 4
    clean_env.setenv("OMPI_COMM_WORLD_RANK", "3")
    clean_env.setenv("OMPI_COMM_WORLD_SIZE", "6")
    clean_env.setenv("OMPI_COMM_WORLD_LOCAL_RANK", "1")
    clean_env.setenv("OMPI_COMM_WORLD_LOCAL_SIZE", "2")
    assert envs.get_rank() == 3 and envs.get_world_size() == 6
    assert envs.get_local_rank() == 1 and envs.get_local_size() == 2
    clean_env.setenv("RANK", "1")            # torchrun / in-tree launcher wins
    clean_env.setenv("WORLD_SIZE", "2")
    clean_env.setenv("LOCAL_RANK", "0")
    clean_env.setenv("LOCAL_WORLD_SIZE", "2")
    assert envs.get_rank() == 1 and envs.get_world_size() == 2
    assert envs.get_local_rank() == 0 and envs.get_local_size() == 2


def test_master_port(clean_env):
    clean_env.setenv("MASTER_PORT", "29500")
    assert envs.get_master_port() == 29601         # never the torchrun agent's own port
    clean_env.setenv("DDLB_MASTER_PORT", "4000")
    assert envs.get_master_port() == 4000
    assert envs.rendezvous_port(3) == 4003
    assert envs.rendezvous_port(199) == 4000
    clean_env.setenv("DDLB_MASTER_ADDR", "10.0.0.1")
    assert envs.get_master_addr() == "10.0.0.1"


def test_options_manager():
    om = OptionsManager({"backend": "nccl", "s": 2, "implementation": None},
                        {"backend": ["nccl", "gloo"], "s": (1, 8)})
    om.parse({"backend": "gloo", "implementation": "pytorch"})
    assert om["backend"] == "gloo" and om.get("s") == 2 and om.get("missing", 7) == 7
    with pytest.raises(ValueError, match="Unknown options"):
        OptionsManager({"a": 1}).parse({"b": 2})
    with pytest.raises(ValueError, match="Must be one of"):
        OptionsManager({"backend": "nccl"}, {"backend": ["nccl"]}).parse({"backend": "mpi"})
    with pytest.raises(ValueError, match="between 1 and 8"):
        OptionsManager({"s": 2}, {"s": (1, 8)}).parse({"s": 9})
    with pytest.raises(ValueError):
        OptionsManager({"s": 2}, {"s": (1, 8)}).parse({"s": "4"})
    with pytest.raises(ValueError):
        OptionsManager({"s": 2}, {"s": (1, 8)}).parse({"s": True})


def test_env_guard(monkeypatch):
    monkeypatch.setenv("DDLB_T_A", "old")
    monkeypatch.delenv("DDLB_T_B", raising=False)
    g = EnvVarGuard({"DDLB_T_A": "new", "DDLB_T_B": "x"})
    assert os.environ["DDLB_T_A"] == "new" and os.environ["DDLB_T_B"] == "x"
    g.close()
    g.close()
    assert os.environ["DDLB_T_A"] == "old" and "DDLB_T_B" not in os.environ
    with EnvVarGuard({"DDLB_T_B": "y"}):
        assert os.environ["DDLB_T_B"] == "y"
    assert "DDLB_T_B" not in os.environ


def test_ucc_env():
    assert setup_ucc_env_vars("nccl") == {}
    assert setup_ucc_env_vars("ucc/tl/cuda") == {"UCC_CL_BASIC_TLS": "cuda"}
    env = setup_ucc_env_vars("ucc/tl/ucp")
    assert env["UCC_CL_BASIC_TLS"] == "ucp" and env["UCX_RNDV_THRESH"] == "0"


def test_lazy_import_does_not_load_torch_cuda_or_native():
    import subprocess
    import sys
    code = ("import sys, ddlb_b200, ddlb_b200.cli, ddlb_b200.primitives.TPColumnwise as c;"
            "import ddlb_b200._native as n;"
            "assert 'torch' not in sys.modules, 'torch imported eagerly';"
            "assert not n.loaded();"
            "c.FusedTPColumnwise; c.PyTorchTPColumnwise; c.FuserTPColumnwise;"
            "assert 'torch' not in sys.modules;"
            "print(ddlb_b200.__version__)")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True,
                         cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    assert out.returncode == 0, out.stderr
    assert out.stdout.strip() == "0.2.0"
