"""Native SMO solvers: the C++ host loop reproduces the tensor restatement exactly; the one-launch CUDA kernel
(csrc/svm.cu) reproduces the host solver (same working-set rule; compared by KKT gap and objective)."""
import pytest
import torch

from harp_b200.daal import classify as C


def _problem(n=400, d=6, seed=0, noise=0.4):
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(n, d, generator=g, dtype=torch.float64)
    w = torch.randn(d, generator=g, dtype=torch.float64)
    y = torch.where(X @ w + noise * torch.randn(n, generator=g, dtype=torch.float64) > 0, 1.0, -1.0).double()
    return X, y


@pytest.mark.parametrize("kernel", ["linear", "rbf"])
def test_host_smo_equals_the_tensor_loop(native_built, kernel):
    X, y = _problem()
    K = (C.kernel_linear(X) if kernel == "linear" else C.kernel_rbf(X, None, 1.5)).contiguous()
    a1, g1 = C._smo_torch(K, y, 1.0, 1e-3, 20000)
    a2, g2 = C._smo_solve(K, y, 1.0, 1e-3, 20000)
    torch.testing.assert_close(a1, a2, rtol=0, atol=1e-12)
    torch.testing.assert_close(g1, g2, rtol=0, atol=1e-10)
    assert abs(float((a2 * y).sum())) < 1e-9 and float(a2.min()) >= 0 and float(a2.max()) <= 1.0 + 1e-12
    m = C.svm_train_binary(X, y, 1.0, kernel, 1.5)
    assert float(((m.decision(X) >= 0) == (y > 0)).double().mean()) > 0.85


@pytest.mark.gpu
@pytest.mark.parametrize("n,kernel", [(400, "linear"), (1500, "rbf"), (37, "linear"), (2500, "linear")])
def test_device_smo_kernel_equals_the_host_solver(native_built, n, kernel):
    X, y = _problem(n=n, seed=n)
    K = (C.kernel_linear(X) if kernel == "linear" else C.kernel_rbf(X, None, 1.5)).contiguous()
    a_h, g_h = C._smo_solve(K, y, 1.0, 1e-3, 50000)
    a_d, g_d = C._smo_solve(K.cuda(), y.cuda(), 1.0, 1e-3, 50000)
    torch.cuda.synchronize()
    # fused multiply-adds on the device round differently, so near-ties pick other pairs and the two runs stop at
    # slightly different points of the same tol-optimal face: compare feasibility, the KKT gap and the objective
    a_d, g_d = a_d.cpu(), g_d.cpu()
    assert float(a_d.min()) >= 0 and float(a_d.max()) <= 1.0 + 1e-12 and abs(float((a_d * y).sum())) < 1e-8
    Q = (y[:, None] * y[None, :]) * K
    torch.testing.assert_close(g_d, Q @ a_d - 1.0, rtol=0, atol=1e-7)               # the maintained gradient is exact
    f = -y * g_d
    up = ((y > 0) & (a_d < 1.0)) | ((y < 0) & (a_d > 0))
    lo = ((y > 0) & (a_d > 0)) | ((y < 0) & (a_d < 1.0))
    assert float(f[up].max() - f[lo].min()) < 1e-3 + 1e-9                           # stopped at a tol-KKT point
    obj = lambda a: float(0.5 * a @ Q @ a - a.sum())
    assert abs(obj(a_d) - obj(a_h)) <= 1e-4 * max(1.0, abs(obj(a_h)))
    torch.testing.assert_close(a_d, a_h, rtol=0, atol=2e-2)
    m = C.svm_train_multiclass(X.cuda(), (y > 0).long().cuda(), 1.0, kernel, 1.5)
    assert float((m.predict(X.cuda()).cpu() == (y > 0).long()).double().mean()) > 0.85
