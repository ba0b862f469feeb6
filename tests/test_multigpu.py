"""Peer-memory gradient all-reduce and the prepare_model data-parallel path on real peers (>= 2 GPUs on one box)."""
import os
import subprocess
import sys

import pytest
import torch


def _why(r):
    """The lines of a failed multi-process run that say what went wrong (the tail alone is torchrun's own traceback)."""
    lines = (r.stdout + "\n" + r.stderr).splitlines()
    keep = [l for l in lines if any(k in l for k in ("Error", "error", "assert", "File \"/", "mismatch", "diverged", "differ"))
            and "torch/distributed" not in l and "ChildFailedError" not in l]
    return "\n".join(keep[-40:]) + "\n--- tail ---\n" + r.stderr[-1500:]

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_peer_memory_ddp(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29600 + world), os.path.join(HERE, "mp_peer_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MP_OK" in r.stdout, _why(r)


@pytest.mark.parametrize("world", [2, 4, 8])
def test_comm_hook_on_stock_ddp(world):
    """b200_train.peer_allreduce_hook behind torch DDP's own register_comm_hook == the default all-reduce."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(29700 + world), os.path.join(HERE, "mp_hook_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "HOOK_OK" in r.stdout, _why(r)
