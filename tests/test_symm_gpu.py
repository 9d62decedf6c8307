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


@pytest.mark.multigpu
@pytest.mark.parametrize("op", ["allreduce", "regroup-allgather", "broadcast-reduce", "push-pull"])
def test_contrib_kmeans_variants_run_on_the_nvlink_kernels(native_built, tmp_path, op):
    """The four tutorial variants with the centroid table as a dense slab: every collective of every iteration is a
    launch of the NVLink kernels (counted), and the result equals the sparse partition-object path."""
    import re
    from harp_b200.io import formats as F
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    out = {}
    for tables in ("dense", "sparse"):
        work = tmp_path / f"{op}-{tables}"
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
               "--master-port", "29537", "-m", "harp_b200.launch.kmeans_contrib", "4000", "12", "16", "2", "8", str(work),
               str(tmp_path / "local"), op, f"--tables={tables}"]
        res = subprocess.run(cmd, capture_output=True, text=True, timeout=150, cwd=str(ROOT),
                             env=dict(os.environ, HB_WAIT_TIMEOUT_S="20"))
        sys.stdout.write(res.stdout[-1500:]); sys.stderr.write(res.stderr[-3000:])
        assert res.returncode == 0
        m = re.search(r"collective kernel launches: (\d+)", res.stdout)
        out[tables] = (F.read_evaluation(str(work / "evaluation"))[1], int(m.group(1)) if m else None)
    assert out["dense"][1] is not None and out["dense"][1] >= 8, out
    assert out["sparse"][1] is None
    assert abs(out["dense"][0] - out["sparse"][0]) <= 1e-3 * abs(out["sparse"][0]), out


@pytest.mark.multigpu
def test_round2_app_kernels_multi_gpu(native_built, tmp_path):
    """PageRank (dense push), subgraph counting (exchange / rotation), WDA-MDS and MF-CCD mappers on 2 GPUs."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29538", str(ROOT / "tests" / "mp_apps_gpu_worker.py"), str(tmp_path)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=dict(os.environ, HB_WAIT_TIMEOUT_S="30"))
    sys.stdout.write(res.stdout[-2000:]); sys.stderr.write(res.stderr[-3000:])
    assert res.returncode == 0 and "APPS_GPU_OK" in res.stdout
