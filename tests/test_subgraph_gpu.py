"""GPU check of the colour-coding DP kernel (csrc/subgraph.cu) against the fp64 torch restatement."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("k,edges", [(5, [(0, 1), (1, 2), (1, 3), (3, 4)]),
                                     (7, [(0, 1), (0, 2), (2, 3), (2, 4), (4, 5), (4, 6)]),
                                     (10, [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (0, 6), (6, 7), (7, 8), (8, 9)])])
def test_colour_coding_pass_device_matches_cpu(native_built, k, edges):
    from harp_b200 import _native
    from harp_b200.apps.graph import SubTemplate, colour_coding_pass
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    g = torch.Generator().manual_seed(k)
    n, deg = 20_000, 6
    src = torch.arange(n).repeat_interleave(deg)
    dst = torch.randint(0, n, (n * deg,), generator=g)
    a = torch.cat([src, dst]); b = torch.cat([dst, src])              # symmetric
    key = torch.unique(a * n + b)
    a, b = key // n, key % n
    keep = a != b
    a, b = a[keep], b[keep]
    ptr = torch.zeros(n + 1, dtype=torch.int64)
    ptr[1:] = torch.cumsum(torch.bincount(a, minlength=n), 0)
    idx = b.to(torch.int32)
    adj_t = {i: [] for i in range(k)}
    for x, y in edges:
        adj_t[x].append(y); adj_t[y].append(x)
    tmpl = SubTemplate(0, frozenset(range(k)), adj_t)
    colours = torch.randint(0, k, (n,), generator=g)
    want = colour_coding_pass(ptr, idx, colours, tmpl, k)[:, 0]
    l0 = _native.launches()
    got = colour_coding_pass(ptr.cuda(), idx.cuda(), colours.cuda(), tmpl, k)[:, 0]
    assert _native.launches() - l0 == k - 1                              # one fused launch per DP step (template edge)
    torch.testing.assert_close(got.double().cpu(), want, rtol=2e-5, atol=1e-3)
    assert float(want.sum()) > 0


def test_pagerank_push_device_matches_cpu(native_built):
    from harp_b200.ops import graphops as GO
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    g = torch.Generator().manual_seed(3)
    n, m = 50_000, 20_000
    deg = torch.randint(0, 12, (m,), generator=g)
    deg[::7] = 0                                              # dangling pages
    ptr = torch.zeros(m + 1, dtype=torch.int64); ptr[1:] = torch.cumsum(deg, 0)
    idx = torch.randint(0, n, (int(ptr[-1]),), generator=g).to(torch.int32)
    src = torch.randperm(n, generator=g)[:m]
    pr = torch.rand(n, generator=g, dtype=torch.float64)
    oc, dc = torch.zeros(n, dtype=torch.float64), torch.zeros(1, dtype=torch.float64)
    GO.pagerank_push(ptr, idx, src, pr, oc, dc)
    og, dg = torch.zeros(n, dtype=torch.float64, device="cuda"), torch.zeros(1, dtype=torch.float64, device="cuda")
    GO.pagerank_push(ptr.cuda(), idx.cuda(), src.cuda(), pr.cuda(), og, dg)
    torch.testing.assert_close(og.cpu(), oc, rtol=1e-12, atol=1e-12)
    torch.testing.assert_close(dg.cpu(), dc, rtol=1e-12, atol=1e-12)
