"""Multi-process (gloo, CPU) tests: collective semantics and the K-means apps end to end."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
# fp64 CPU paths agree to rounding; one rank per GPU (HARP_MP_TESTS_GPU=1) runs fp32 kernels with atomically ordered sums
TOL = 1e-5 if os.environ.get("HARP_MP_TESTS_GPU", "0") == "1" else 1e-9
_PORT = [29700]


def _torchrun(nproc, args, timeout=600):
    _PORT[0] += 1
    # HARP_MP_TESTS_GPU=1 runs the same launchers one rank per GPU over NCCL (a multi-GPU box; tolerances that assume the
    # fp64 CPU paths then apply to fp32 / TF32 device kernels, so this is a smoke run of the device routes, not the gate)
    env = dict(os.environ, OMP_NUM_THREADS="2")
    if os.environ.get("HARP_MP_TESTS_GPU", "0") != "1":
        env["HARP_FORCE_CPU"] = "1"
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_PORT[0])] + args
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=str(ROOT))
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    return res.stdout


@pytest.mark.parametrize("nproc", [2, 4])
def test_collectives_verify(nproc):
    out = _torchrun(nproc, [str(ROOT / "tests" / "mp_collectives_worker.py")])
    assert f"COLLECTIVES_OK {nproc}" in out


def test_kmeans_regroup_allgather_matches_oracle(tmp_path):
    """README config (1000 pts, 10 centroids, dim 100, 2 workers, 10 iters) vs the float64 oracle."""
    from harp_b200.io import formats as F
    from oracle import kmeans as oracle
    work = tmp_path / "work"
    _torchrun(2, ["-m", "harp_b200.launch.kmeans", "1000", "10", "100", "5", "2", "2", "10", str(work), str(tmp_path / "local")])
    got = F.read_centroids_text(str(work / "centroids" / "out" / "output"), 100)
    files = sorted((work / "data").iterdir())
    assert len(files) == 10
    pts = np.concatenate([F.read_kmeans_points_bin(str(f), 100) for f in files])
    init = F.read_centroids_text(str(work / "centroids" / "init_centroids"), 100)
    want, hist = oracle.run(pts, init, 10)
    assert got.shape == (10, 100)
    np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-9)
    assert all(a >= b - 1e-6 for a, b in zip(hist, hist[1:]))          # Lloyd is monotone


@pytest.mark.parametrize("op", ["allreduce", "regroup-allgather", "broadcast-reduce", "push-pull"])
def test_contrib_kmeans_quality_gate(tmp_path, op):
    """contrib/test_scripts/km.sh: 1000 x 10, 10 centroids, 2 mappers, 100 iterations -> 7.0 < MSE < 7.8."""
    from harp_b200.io import formats as F
    work = tmp_path / op
    _torchrun(2, ["-m", "harp_b200.launch.kmeans_contrib", "1000", "10", "10", "2", "100", str(work),
                  str(tmp_path / "local"), op])
    name, mse = F.read_evaluation(str(work / "evaluation"))
    assert name == "MSE" and 7.0 < mse < 7.8, mse
    assert (work / "out" / "part-m-00000").exists()


def test_kmeans_rotation_matches_regroup_allgather(tmp_path):
    from harp_b200.io import formats as F
    a, b = tmp_path / "a", tmp_path / "b"
    _torchrun(2, ["-m", "harp_b200.launch.kmeans", "600", "8", "12", "2", "2", "1", "5", str(a), str(tmp_path / "l")])
    _torchrun(2, ["-m", "harp_b200.launch.kmeans_rotation", "600", "8", "12", "2", "2", "1", "5", str(b), str(tmp_path / "l")])
    ca = F.read_centroids_text(str(a / "centroids" / "out" / "output"), 12)
    cb = F.read_centroids_text(str(b / "centroids" / "out" / "output"), 12)
    np.testing.assert_allclose(ca, cb, rtol=TOL, atol=TOL)


def test_kmeans_launcher_device_engine_matches_the_table_path(tmp_path):
    """--engine=device runs the iterations in KMeansEngine (sums + allreduce here on CPU, the fused NVLink kernel on
    GPUs) instead of the table regroup/allgather; same centroids."""
    from harp_b200.io import formats as F
    a, b = tmp_path / "a", tmp_path / "b"
    args = ["600", "8", "12", "2", "2", "1", "6"]
    _torchrun(2, ["-m", "harp_b200.launch.kmeans"] + args + [str(a), str(tmp_path / "l")])
    _torchrun(2, ["-m", "harp_b200.launch.kmeans"] + args + [str(b), str(tmp_path / "l"), "--engine=device"])
    ca = F.read_centroids_text(str(a / "centroids" / "out" / "output"), 12)
    cb = F.read_centroids_text(str(b / "centroids" / "out" / "output"), 12)
    np.testing.assert_allclose(ca, cb, rtol=TOL, atol=TOL)


def test_kmeans_resume_continues_from_a_centroid_dump(tmp_path):
    """6 iterations, then --resume for 4 more == 10 iterations in one go (same data, same initial centroids)."""
    from harp_b200.io import formats as F
    a, b = tmp_path / "a", tmp_path / "b"
    args = ["600", "8", "12", "2", "2", "1"]
    _torchrun(2, ["-m", "harp_b200.launch.kmeans"] + args + ["10", str(a), str(tmp_path / "l")])
    _torchrun(2, ["-m", "harp_b200.launch.kmeans"] + args + ["6", str(b), str(tmp_path / "l")])
    _torchrun(2, ["-m", "harp_b200.launch.kmeans"] + args + ["4", str(b), str(tmp_path / "l"),
                  f"--resume={b / 'centroids' / 'out' / 'output'}"])
    ca = F.read_centroids_text(str(a / "centroids" / "out" / "output"), 12)
    cb = F.read_centroids_text(str(b / "centroids" / "out" / "output"), 12)
    np.testing.assert_allclose(ca, cb, rtol=TOL, atol=TOL)


def test_numpy_session_api():
    """python/harp_session.py of the reference (a py4j stub there): barrier / broadcast / allreduce / reduce /
    allgather on NumPy arrays, 1 and 3 workers."""
    out = _torchrun(3, [str(ROOT / "tests" / "mp_session_worker.py")])
    assert "SESSION_OK 3" in out
    res = subprocess.run([sys.executable, str(ROOT / "tests" / "mp_session_worker.py")], capture_output=True, text=True,
                         timeout=300, cwd=str(ROOT), env=dict(os.environ, PYTHONPATH=str(ROOT)))
    assert res.returncode == 0 and "SESSION_OK 1" in res.stdout, res.stdout + res.stderr[-2000:]


def test_secondary_apps_end_to_end(tmp_path):
    """PageRank, iterative SVM, MLR, LDA-CVB (static + push/pull), NN, random forest, subgraph counting
    (allgather + SAHAD rotation), WDA-MDS and the SGX-cost-model K-means on generated inputs, 2 workers."""
    out = _torchrun(2, [str(ROOT / "tests" / "mp_apps_worker.py"), str(tmp_path)], timeout=900)
    assert "APPS_OK 2" in out


def test_mf_sgd_and_ccd_launchers(tmp_path):
    from harp_b200.io import datagen, formats as F
    r, c, v = datagen.synth_ratings(300, 120, 9000, rank=4, seed=1, noise=0.05)
    (tmp_path / "train").mkdir(); (tmp_path / "test").mkdir()
    idx = np.arange(9000)
    for k, part in enumerate(np.array_split(idx[:8000], 4)):
        F.write_coo(str(tmp_path / "train" / f"part_{k}"), r[part], c[part], v[part])
    F.write_coo(str(tmp_path / "test" / "part_0"), r[8000:], c[8000:], v[8000:])
    out = _torchrun(2, ["-m", "harp_b200.launch.sgd", str(tmp_path / "train"), "8", "0.05", "0.02", "25", "100", "2", "1", "2",
                        "2000", str(tmp_path / "work"), str(tmp_path / "test")])
    rmse = float(open(tmp_path / "work" / "model" / "evaluation").read())
    assert rmse < 0.5 * float(np.std(v)), rmse
    ids, W = F.read_mf_model(str(tmp_path / "work" / "model" / "W-0"))
    assert W.shape[1] == 8 and len(ids) > 50 and (tmp_path / "work" / "model" / "H-1").exists()
    out = _torchrun(2, ["-m", "harp_b200.launch.ccd", str(tmp_path / "train"), "8", "0.05", "6", "2", "1", "2000",
                        str(tmp_path / "workc"), str(tmp_path / "test")])
    rmse_c = float(open(tmp_path / "workc" / "model" / "evaluation").read())
    assert rmse_c < 0.7 * float(np.std(v)), rmse_c


def test_lda_launcher(tmp_path):
    from harp_b200.io import datagen
    datagen.generate_corpus_files(60, 50, 3, 25, 4, str(tmp_path / "docs"), seed=2)
    out = _torchrun(2, ["-m", "harp_b200.launch.lda", str(tmp_path / "docs"), "3", "0.1", "0.01", "8", "100", "100", "2", "1", "2",
                        "2000", str(tmp_path / "work"), "true", "--print_interval=4"])
    ll = float(open(tmp_path / "work" / "model" / "evaluation").read())
    assert ll < 0 and (tmp_path / "work" / "model" / "tmp_word_model" / "8" / "0").exists()


def test_daal_launcher_matches_numpy(tmp_path):
    from harp_b200.io import datagen
    datagen.generate_dense_csv(400, 8, 4, str(tmp_path / "cov"), seed=0)
    out = _torchrun(2, ["-m", "harp_b200.launch.daal", "cov", "2", "2", "2000", "1", str(tmp_path / "cov"), str(tmp_path / "work"), "8"])
    assert "Harp-DAAL cov finished" in out and "cov.covariance:" in out
