"""Multi-GPU check of the NVLink peer-memory collectives (needs >= 2 GPUs on the box)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = [pytest.mark.gpu]
ROOT = Path(__file__).resolve().parent.parent


@pytest.mark.multigpu
@pytest.mark.parametrize("backend", ["auto"])
def test_symm_collectives_multi_gpu(native_built, backend):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else (4 if n < 8 else 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           str(ROOT / "tests" / "mp_symm_worker.py"), backend]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    sys.stdout.write(res.stdout[-4000:]); sys.stderr.write(res.stderr[-4000:])
    assert res.returncode == 0 and "SYMM_OK" in res.stdout


@pytest.mark.multigpu
def test_rotation_engines_multi_gpu(native_built):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29534", str(ROOT / "tests" / "mp_engines_worker.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=400)
    sys.stdout.write(res.stdout[-3000:]); sys.stderr.write(res.stderr[-3000:])
    assert res.returncode == 0 and "ENGINES_OK" in res.stdout


@pytest.mark.parametrize("worker,token,port", [("mp_symm_worker.py", "SYMM_OK", 29535), ("mp_engines_worker.py", "ENGINES_OK", 29536)])
def test_workers_in_a_world_of_one(native_built, worker, token, port):
    """The same workers on ONE GPU: every peer pointer is the rank's own slab, so flag protocols, ticket reuse, the
    row kernels and the engines' collectives all run on a single-GPU box (what the round-end driver has)."""
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=1",
           "--master-addr", "127.0.0.1", "--master-port", str(port), str(ROOT / "tests" / worker)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=400)
    sys.stdout.write(res.stdout[-4000:]); sys.stderr.write(res.stderr[-4000:])
    assert res.returncode == 0 and token in res.stdout
