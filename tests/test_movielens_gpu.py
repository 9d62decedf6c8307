"""The reference's own MF-SGD acceptance gate (``ml/java/test_scripts/mfsgd.sh:62-75``): MovieLens, r = 40,
lambda = 0.05, epsilon = 0.002, 200 iterations -> 0.80 < test RMSE < 0.84 (the script's comment: 0.8345).

The data set ships with the reference (``datasets/daal_als/movielens-{train,test}``); the test looks for it in
``$HARP_MOVIELENS_DIR``, in ``data_cache/movielens`` of this repo and in the reference tree, and skips when none
is there (nothing is downloaded)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _dataset():
    for cand in (os.environ.get("HARP_MOVIELENS_DIR"), ROOT / "data_cache" / "movielens",
                 "/root/reference/datasets/daal_als"):
        if cand and (Path(cand) / "movielens-train").is_dir() and (Path(cand) / "movielens-test").is_dir():
            return Path(cand)
    return None


def test_movielens_test_rmse_is_inside_the_reference_window(native_built, tmp_path):
    data = _dataset()
    if data is None:
        pytest.skip("MovieLens files of the reference are not available on this machine")
    work = tmp_path / "work"
    cmd = [sys.executable, "-m", "harp_b200.launch.sgd", str(data / "movielens-train"), "40", "0.05", "0.002", "200",
           "100", "1", "16", "2", "20000", str(work), str(data / "movielens-test"), "--rmse_interval=50"]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=1200, cwd=str(ROOT),
                         env=dict(os.environ, HARP_LOG="INFO"))
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    rmse = float((work / "model" / "evaluation").read_text().split()[0])
    print(res.stdout[-300:], "\n".join(l for l in res.stderr.splitlines() if "RMSE" in l)[-600:])
    assert 0.80 < rmse < 0.84, rmse
    assert (work / "model" / "W-0").exists() and (work / "model" / "H-0").exists()
