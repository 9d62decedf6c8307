"""GPU numerics of the linear-algebra kernels (csrc/linalg.cu) against fp64 references."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def _norm_sign(R):
    s = torch.sign(R.diag()); s[s == 0] = 1
    return R * s[:, None]


@pytest.mark.parametrize("n,d,view", [(100_037, 64, False), (300_000, 32, False), (70_001, 16, False), (50_000, 20, False),
                                       (9_000, 5, False), (40_000, 64, True), (300, 64, False), (1, 8, False)])
def test_tsqr_matches_fp64_qr(n, d, view):
    from harp_b200.ops import linalg as KL
    dev = _dev()
    g = torch.Generator(device="cpu").manual_seed(n + d)
    base = torch.randn(n, d + (8 if view else 0), generator=g).to(dev)
    X = base[:, 4:4 + d] if view else base            # a strided view exercises the non-bulk staging path
    assert KL.tsqr_supported(X)
    R = _norm_sign(KL.tsqr_r(X).double())
    assert float(R.tril(-1).abs().max()) == 0.0
    X64 = X.double()
    G = X64.t() @ X64
    assert float((R.t() @ R - G).abs().max() / G.abs().max()) < 2e-5
    if n >= d:
        R64 = _norm_sign(torch.linalg.qr(X64, mode="r").R)
        assert float((R - R64).abs().max() / R64.abs().max()) < 1e-4


def test_tsqr_keeps_orthogonality_where_the_gram_route_loses_it():
    """Columns spanning five orders of magnitude with near-dependencies (cond ~ 1e5): Q = X R^-1 from the Householder
    triangle stays orthogonal to fp32 level; a Cholesky of the fp32 Gram matrix squares the condition number."""
    from harp_b200.ops import linalg as KL
    dev = _dev()
    g = torch.Generator(device="cpu").manual_seed(5)
    n, d = 200_000, 32
    U = torch.linalg.qr(torch.randn(n, d, generator=g, dtype=torch.float64)).Q
    V = torch.linalg.qr(torch.randn(d, d, generator=g, dtype=torch.float64)).Q
    sv = torch.logspace(0, -5, d, dtype=torch.float64)
    X = ((U * sv) @ V.t()).float().to(dev)
    R = KL.tsqr_r(X).double()
    Q = X.double() @ torch.linalg.inv(R)
    err = float((Q.t() @ Q - torch.eye(d, device=dev, dtype=torch.float64)).abs().max())
    assert err < 5e-2, err
    sv_est = torch.linalg.svdvals(R)
    assert float(((sv_est.cpu() - sv) / sv).abs().max()) < 5e-2     # even the smallest singular value survives


@pytest.mark.parametrize("f,batch", [(10, 1000), (32, 513), (64, 200), (3, 7)])
def test_batched_cholesky_factor_and_solve(f, batch):
    from harp_b200.ops import linalg as KL
    dev = _dev()
    g = torch.Generator(device="cpu").manual_seed(f)
    M = torch.randn(batch, f, f + 5, generator=g, dtype=torch.float64)
    A = (M @ M.transpose(1, 2) + 0.5 * torch.eye(f, dtype=torch.float64)).to(dev)
    B = torch.randn(batch, f, generator=g, dtype=torch.float64).to(dev)
    X, L, bad = KL.potrs_batched(A.float(), B.float(), want_factor=True)
    assert bad == 0
    L64 = torch.linalg.cholesky(A)
    X64 = torch.cholesky_solve(B.unsqueeze(-1), L64).squeeze(-1)
    assert float((L.double() - L64).abs().max() / L64.abs().max()) < 1e-4
    assert float((X.double() - X64).abs().max() / X64.abs().max()) < 2e-3
    # a matrix that is not positive definite is counted, not silently "solved"
    A2 = A.float().clone(); A2[0] = -A2[0]
    _, _, bad = KL.potrs_batched(A2, B.float())
    assert bad == 1


@pytest.mark.parametrize("f", [8, 10, 32, 50, 64])
def test_als_solve_matches_the_fp64_restatement(f):
    from harp_b200.daal import models as M
    from harp_b200.ops import linalg as KL
    dev = _dev()
    rng = np.random.default_rng(f)
    n_u, n_i, nnz = 700, 400, 30_000
    key = np.unique(rng.integers(0, n_u * n_i, nnz))
    r, c = torch.from_numpy(key // n_i), torch.from_numpy(key % n_i)
    v = torch.from_numpy(rng.integers(1, 6, key.size).astype(np.float64))
    r[r == 3] = 4                                                       # user 3 has no ratings at all
    ptr, idx, val = M._csr_from_coo(r, c, v, n_u)
    Y = torch.from_numpy(rng.random((n_i, f)) * 0.5)
    want = M._als_solve_side(ptr, idx, val, Y, n_u, 0.1, 20.0)
    F = KL.als_pad(f)
    Yf = torch.zeros(n_i, F, dtype=torch.float32, device=dev); Yf[:, :f] = Y.float().to(dev)
    out = torch.full((n_u, F), 7.0, dtype=torch.float32, device=dev)
    side = KL.ALSSide(ptr.to(dev), idx.to(dev), val.to(dev))
    KL.als_solve(side, Yf, out, f, 0.1, 20.0)
    assert int(side.info.item()) == 0
    got = out.cpu().double()
    assert float(got[:, f:].abs().max()) == 0.0 if F > f else True
    assert float(got[3].abs().max()) == 0.0
    assert float((got[:, :f] - want).abs().max() / want.abs().max()) < 2e-3
    # a subset of rows (what a rank owns in the distributed protocol) leaves the others untouched
    rows = torch.arange(1, n_u, 2, device=dev)
    out2 = torch.full((n_u, F), 7.0, dtype=torch.float32, device=dev)
    KL.als_solve(KL.ALSSide(ptr.to(dev), idx.to(dev), val.to(dev), rows), Yf, out2, f, 0.1, 20.0)
    assert torch.equal(out2[1::2], out[1::2]) and float((out2[0::2] - 7.0).abs().max()) == 0.0


def test_als_implicit_device_route_tracks_the_cpu_route():
    from harp_b200.daal import models as M
    dev = _dev()
    rng = np.random.default_rng(0)
    U, I, f = 300, 200, 10
    key = np.unique(rng.integers(0, U * I, 9000))
    r, c = torch.from_numpy(key // I), torch.from_numpy(key % I)
    v = torch.ones(key.size, dtype=torch.float64)
    Xc, Yc = M.als_implicit(r, c, v, U, I, f, 5, 0.1, 20.0)
    Xg, Yg = M.als_implicit(r.to(dev), c.to(dev), v.to(dev), U, I, f, 5, 0.1, 20.0)
    pc, pg = Xc @ Yc.t(), (Xg @ Yg.t()).cpu()
    assert float((pc - pg).abs().max()) < 5e-3


def test_daal_qr_svd_pca_run_on_the_kernel():
    from harp_b200 import _native
    from harp_b200.daal import linalg as L
    dev = _dev()
    g = torch.Generator(device="cpu").manual_seed(1)
    X = (torch.randn(60_000, 24, generator=g) * torch.linspace(1, 3, 24)).to(dev)
    before = _native.launches()
    Q, R = L.qr_distributed(X)
    assert _native.launches() > before
    assert float((Q.double() @ R - X.double()).abs().max()) < 1e-3
    assert float((Q.t() @ Q - torch.eye(24, device=dev)).abs().max()) < 1e-3
    U, S, Vh = L.svd_distributed(X)
    want = torch.linalg.svdvals(X.double())
    assert float(((S - want) / want).abs().max()) < 1e-4
    res = L.pca_svd(X)
    ev = torch.linalg.eigvalsh(torch.corrcoef(X.double().t())).flip(0)
    assert float((res.eigenvalues - ev).abs().max() / ev.max()) < 1e-3
