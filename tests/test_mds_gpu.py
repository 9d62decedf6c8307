"""GPU check of the fused WDA-MDS pass (csrc/mds.cu) against the fp64 torch restatement."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,m,dim,shift", [(1000, 1000, 3, 0.0), (3001, 517, 2, 0.3), (4096, 64, 8, 0.05)])
def test_mds_bc_stress_matches_fp64(native_built, n, m, dim, shift):
    from harp_b200.ops import mds as KM
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    g = torch.Generator().manual_seed(n + dim)
    P = torch.rand(n, 4, generator=g, dtype=torch.float64)
    ids = torch.randperm(n, generator=g)[:m].sort().values
    D = (P[ids][:, None, :] - P[None, :, :]).pow(2).sum(2).sqrt()
    X = torch.rand(n, dim, generator=g, dtype=torch.float64)
    X[5] = X[ids[0]]                                         # a coincident pair: the ratio is defined as 0 there
    BXc, stc = KM.bc_stress(D, ids, X, shift)
    BXg, stg = KM.bc_stress(D.cuda(), ids.cuda(), X.cuda(), shift)
    torch.testing.assert_close(BXg.double().cpu(), BXc, rtol=2e-4, atol=2e-4 * float(BXc.abs().max()))
    torch.testing.assert_close(stg.cpu(), stc, rtol=1e-4, atol=1e-6)
