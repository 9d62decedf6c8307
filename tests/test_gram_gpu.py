"""GPU numerics of the Gram / moments kernels vs a float64 PyTorch reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,d", [(1000, 256), (5000, 512), (777, 32), (4096, 1024), (20000, 96), (100, 320)])
def test_gram_tc(native_built, n, d):
    from harp_b200.ops.gram import gram
    X = torch.randn(n, d, device="cuda")
    G = gram(X)
    torch.cuda.synchronize()
    ref = (X.double().t() @ X.double())
    err = (G.double() - ref).abs().max().item()
    scale = ref.abs().max().item()
    assert err <= 2e-3 * scale, (err, scale)                  # TF32 products
    assert torch.equal(G, G.t())                               # lower triangle is a copy of the upper one
    G2 = gram(X, out=G.clone(), accumulate=True)
    torch.cuda.synchronize()
    assert (G2.double() - 2 * ref).abs().max().item() <= 4e-3 * scale


@pytest.mark.parametrize("n,d", [(1000, 100), (513, 37), (3000, 256)])
def test_gram_simt(native_built, n, d):
    from harp_b200.ops.gram import gram
    X = torch.randn(n, d, device="cuda")
    G = gram(X, force_simt=True)
    torch.cuda.synchronize()
    ref = (X.double().t() @ X.double())
    torch.testing.assert_close(G.double(), ref, rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("n,d", [(10001, 77), (10001, 128), (50000, 1024), (37, 260), (3, 4)])
def test_moments(native_built, n, d):
    from harp_b200.ops.gram import moments
    X = torch.randn(n, d, device="cuda") * 3 - 1
    m = moments(X)
    torch.cuda.synchronize()
    torch.testing.assert_close(m.sum, X.double().sum(0), rtol=1e-9, atol=1e-6)
    torch.testing.assert_close(m.sumsq, (X.double() ** 2).sum(0), rtol=1e-9, atol=1e-6)
    assert torch.equal(m.min, X.min(0).values) and torch.equal(m.max, X.max(0).values)


@pytest.mark.parametrize("n,d", [(1000, 256), (5000, 512), (777, 32), (4096, 1024), (100, 320), (0, 64)])
def test_gram_fused_single_gpu(native_built, n, d):
    """Gram + (degenerate, one-rank) allreduce + mirror in one launch == the two-launch path, twice in a row."""
    from harp_b200.ops.gram import FusedGram, gram
    st = torch.empty(FusedGram.state_bytes(d), dtype=torch.uint8, device="cuda")
    fg = FusedGram(st, [st.data_ptr()], 0, d)
    for rep in range(2):
        X = torch.randn(n, d, device="cuda")
        G = fg(X).clone()
        torch.cuda.synchronize()
        ref = (X.double().t() @ X.double())
        scale = max(ref.abs().max().item(), 1.0) if n else 1.0
        assert (G.double() - ref).abs().max().item() <= 2e-3 * scale
        assert torch.equal(G, G.t())
        if n:
            G2 = gram(X)
            torch.cuda.synchronize()
            torch.testing.assert_close(G, G2, rtol=1e-5, atol=1e-4 * scale)   # split-K reductions land in any order


def test_covariance_of_data_with_a_large_mean(native_built):
    """mean 100, spread 1: the two-pass route (exact float64 column sums, then the tensor-core Gram of the centred
    rows) must reproduce the float64 covariance to rtol 1e-4 in the diagonal and 1e-3 of sd_i sd_j off it; the
    one-pass G - N m mᵀ route, forced, is shown NOT to (that is why the default switches)."""
    from harp_b200.daal.linalg import covariance
    g = torch.Generator(device="cuda").manual_seed(0)
    n, d = 200_000, 256
    A = torch.randn(d, d, device="cuda", generator=g) * 0.2 + torch.eye(d, device="cuda")
    X = (torch.randn(n, d, device="cuda", generator=g) @ A + 100.0).contiguous()
    ref = torch.cov(X.double().t())
    c = covariance(X)
    torch.testing.assert_close(c.mean, X.double().mean(0), rtol=1e-8, atol=1e-9)   # fp32 partial sums of four rows, folded in fp64
    sd = ref.diag().sqrt()
    torch.testing.assert_close(c.covariance.diag(), ref.diag(), rtol=1e-4, atol=0)
    assert float(((c.covariance - ref).abs() / torch.outer(sd, sd)).max()) < 1e-3
    bad = covariance(X, center=False)
    assert float(((bad.covariance - ref).abs() / torch.outer(sd, sd)).max()) > 1e-2


def test_xty_and_linear_regression_without_copy(native_built):
    from harp_b200.daal.linalg import linear_regression_normeq
    from harp_b200.ops.gram import xty
    g = torch.Generator(device="cuda").manual_seed(1)
    n, d, k = 100_000, 128, 2
    X = (torch.randn(n, d, device="cuda", generator=g) + 3.0).contiguous()
    beta = torch.randn(k, d + 1, device="cuda", generator=g, dtype=torch.float64)
    Y = (X.double() @ beta[:, 1:].t() + beta[:, 0] + 0.01 * torch.randn(n, k, device="cuda", generator=g, dtype=torch.float64)).float()
    out = xty(X, Y)
    torch.testing.assert_close(out[0], X.double().sum(0), rtol=1e-12, atol=1e-6)
    torch.testing.assert_close(out[1:], Y.double().t() @ X.double(), rtol=1e-10, atol=1e-4)
    peak0 = torch.cuda.max_memory_allocated()
    model = linear_regression_normeq(X, Y)
    assert torch.cuda.max_memory_allocated() - peak0 < X.numel() * 2          # no [1 | X | Y] copy of X
    torch.testing.assert_close(model.beta, beta, rtol=5e-2, atol=5e-2)
    pred = model.predict(X)
    assert float((pred - Y.double()).pow(2).mean().sqrt()) < 0.05


@pytest.mark.gpu
@pytest.mark.parametrize("n,d", [(100_003, 256), (50_001, 1024), (7_777, 288), (15, 64), (300_000, 512), (0, 64)])
def test_gram_fused_with_column_moments(native_built, n, d):
    """The statistics warps of the Gram kernel: sum / sum of squares / min / max of every column from the SAME launch
    that produces the Gram matrix, against fp64 torch; repeated so that re-initialisation of the arrays is covered."""
    from harp_b200.ops.gram import FusedGram
    st = torch.empty(FusedGram.state_bytes(d), dtype=torch.uint8, device="cuda")
    fg = FusedGram(st, [st.data_ptr()], 0, d)
    g = torch.Generator(device="cuda").manual_seed(n + d)
    for rep in range(2):
        X = torch.randn(n, d, device="cuda", generator=g) * 3 + torch.linspace(-50, 50, d, device="cuda")
        G, m = fg(X, want_moments=True)
        torch.cuda.synchronize()
        assert m.n == n
        if n == 0:
            assert float(m.sum.abs().max()) == 0.0 and bool(torch.isinf(m.min).all())
            continue
        Xd = X.double()
        torch.testing.assert_close(m.sum, Xd.sum(0), rtol=1e-6, atol=1e-6 * n)
        torch.testing.assert_close(m.sumsq, (Xd * Xd).sum(0), rtol=1e-6, atol=1e-6 * n)
        assert torch.equal(m.min, X.min(0).values) and torch.equal(m.max, X.max(0).values)
        ref = Xd.t() @ Xd
        assert float((G.double() - ref).abs().max() / ref.abs().max()) < 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("fused_stats", ["0", "1"])
def test_covariance_small_mean_route(native_built, monkeypatch, fused_stats):
    """No centring needed: the decision pass IS the answer (Gram + moments once, no third pass); with
    HB_GRAM_FUSED_STATS=1 the moments come out of the Gram launch itself."""
    from harp_b200 import _native
    from harp_b200.daal import linalg as L
    monkeypatch.setenv("HB_GRAM_FUSED_STATS", fused_stats)
    X = torch.randn(200_000, 256, device="cuda") * 2 + 0.3
    _native.reset_launches()
    res = L.covariance(X)
    # prefix moments that pick the route (2) + Gram with the moments inside (2: init + kernel) or Gram (1) + moments (2)
    assert _native.launches() == (4 if fused_stats == "1" else 5)
    ref = torch.cov(X.double().t())
    assert float((res.covariance - ref).abs().max()) < 2e-2
    torch.testing.assert_close(res.mean, X.double().mean(0), rtol=1e-6, atol=1e-6)
