"""GPU checks of the MF kernels: SGD pass reduces the loss and matches the sequential CPU restatement
statistically; squared error and CCD update match PyTorch references."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _problem(nr=3000, nc=500, nnz=60000, rank=32, seed=0):
    g = torch.Generator().manual_seed(seed)
    Wt = torch.randn(nr, 8, generator=g) / math.sqrt(8)
    Ht = torch.randn(nc, 8, generator=g) / math.sqrt(8)
    r = torch.randint(0, nr, (nnz,), generator=g)
    c = torch.randint(0, nc, (nnz,), generator=g)
    v = (Wt[r] * Ht[c]).sum(1) * 2 + 3 + 0.05 * torch.randn(nnz, generator=g)
    return r, c, v.float()


def test_sq_err_matches_torch(native_built):
    from harp_b200.ops import mf
    r, c, v = _problem()
    W = torch.rand(3000, 32, device="cuda"); H = torch.rand(500, 32, device="cuda")
    got = mf.sq_err(W, H, r.cuda(), c.cuda(), v.cuda())
    ref = ((v.cuda() - (W[r.cuda()] * H[c.cuda()]).sum(1)).double() ** 2).sum()
    torch.testing.assert_close(got[0], ref, rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("rank,B", [(32, 16), (128, 148), (40, 64), (256, 8)])
def test_sgd_pass_trains_everything_and_reduces_loss(native_built, rank, B):
    from harp_b200.ops import mf
    r, c, v = _problem(rank=rank)
    r, c, v = r.cuda(), c.cuda(), v.cuda()
    B = min(B, mf.max_blocks(rank))
    panel = mf.build_panel(r, c, v, B)
    assert panel.nnz == r.numel() and int(panel.blk_seg_ptr[-1]) == panel.seg_col.numel()
    g = torch.Generator(device="cuda").manual_seed(1)
    W = torch.rand(3000, rank, device="cuda", generator=g) / math.sqrt(rank)
    H = torch.rand(500, rank, device="cuda", generator=g) / math.sqrt(rank)
    e0 = float(mf.sq_err(W, H, r, c, v))
    for it in range(15):
        n = mf.sgd_pass(panel, W, H, 0.01, 0.02)
        assert int(n.item()) == r.numel()                       # every rating exactly once per pass
    torch.cuda.synchronize()
    e1 = float(mf.sq_err(W, H, r, c, v))
    assert math.isfinite(e1) and e1 < 0.15 * e0, (e0, e1)
    assert int(panel.col_ticket.min()) == int(panel.col_ticket.max()) == 15 * B


@pytest.mark.parametrize("rank,B,C", [(32, 16, 32), (128, 148, 296), (64, 37, 100)])
def test_sgd_pass_dynamic_schedule(native_built, rank, B, C):
    """Dynamic schedule (any free, not-yet-visited column-block; B row-blocks x C column-blocks): every rating
    exactly once per pass, locks back at rest, loss goes down like the ring schedule's."""
    from harp_b200.ops import mf
    r, c, v = _problem(rank=rank)
    r, c, v = r.cuda(), c.cuda(), v.cuda()
    B = min(B, mf.max_blocks(rank))
    panel = mf.build_panel(r, c, v, B, C=C)
    assert panel.ncolblocks == C and panel.blk_seg_ptr.numel() == B * C + 1
    g = torch.Generator(device="cuda").manual_seed(1)
    W = torch.rand(3000, rank, device="cuda", generator=g) / math.sqrt(rank)
    H = torch.rand(500, rank, device="cuda", generator=g) / math.sqrt(rank)
    e0 = float(mf.sq_err(W, H, r, c, v))
    for it in range(15):
        n = mf.sgd_pass(panel, W, H, 0.01, 0.02)
        assert int(n.item()) == r.numel()
    torch.cuda.synchronize()
    e1 = float(mf.sq_err(W, H, r, c, v))
    assert math.isfinite(e1) and e1 < 0.15 * e0, (e0, e1)
    assert int(panel.col_ticket.abs().max()) == 0


def test_sgd_pass_matches_cpu_restatement(native_built):
    """Same panel, same order where it matters: one CTA (B=1) with ratings in distinct rows per column run
    is fully sequential per column, so the GPU result must match the CPU loop closely."""
    from harp_b200.ops import mf
    g = torch.Generator().manual_seed(3)
    nr, nc, rank = 400, 40, 16
    r = torch.arange(nr).repeat_interleave(1)                   # each row appears once -> no W conflicts
    c = torch.randint(0, nc, (nr,), generator=g)
    v = torch.rand(nr, generator=g) * 4 + 1
    W0 = torch.rand(nr, rank, generator=g) * 0.3
    H0 = torch.rand(nc, rank, generator=g) * 0.3
    panel_cpu = mf.build_panel(r, c, v, 1)
    Wc, Hc = W0.double().clone(), H0.double().clone()
    mf.sgd_pass(panel_cpu, Wc, Hc, 0.05, 0.01)
    panel_gpu = mf.build_panel(r.cuda(), c.cuda(), v.cuda(), 1)
    Wg, Hg = W0.cuda().clone(), H0.cuda().clone()
    mf.sgd_pass(panel_gpu, Wg, Hg, 0.05, 0.01)
    torch.cuda.synchronize()
    torch.testing.assert_close(Wg.cpu().double(), Wc, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(Hg.cpu().double(), Hc, rtol=1e-4, atol=1e-5)


def test_sgd_deadline_bounds_the_pass(native_built):
    from harp_b200.ops import mf
    r, c, v = _problem(nr=200000, nc=20000, nnz=4_000_000, rank=128, seed=5)
    r, c, v = r.cuda(), c.cuda(), v.cuda()
    B = min(148, mf.max_blocks(128))
    panel = mf.build_panel(r, c, v, B)
    W = torch.rand(200000, 128, device="cuda") * 0.1; H = torch.rand(20000, 128, device="cuda") * 0.1
    full = int(mf.sgd_pass(panel, W, H, 0.001, 0.01).item())
    part = int(mf.sgd_pass(panel, W, H, 0.001, 0.01, deadline_ms=0.05).item())
    torch.cuda.synchronize()
    assert full == 4_000_000 and 0 <= part < full


def test_ccd_update_matches_cpu(native_built):
    from harp_b200.ops import mf
    g = torch.Generator().manual_seed(0)
    nr, nc, nnz = 300, 200, 5000
    r = torch.randint(0, nr, (nnz,), generator=g); c = torch.randint(0, nc, (nnz,), generator=g)
    order = torch.argsort(r, stable=True)
    r, c = r[order], c[order]
    ptr = torch.zeros(nr + 1, dtype=torch.int32); ptr[1:] = torch.cumsum(torch.bincount(r, minlength=nr), 0)
    resid = torch.randn(nnz, generator=g)
    Wf = torch.randn(nr, generator=g); Hf = torch.randn(nc, generator=g)
    rc, wc = resid.clone(), Wf.clone()
    mf.ccd_update(ptr, c.int(), rc, None, wc, Hf, 0.1)
    rg, wg = resid.cuda(), Wf.cuda()
    mf.ccd_update(ptr.cuda(), c.int().cuda(), rg, None, wg, Hf.cuda(), 0.1)
    torch.cuda.synchronize()
    torch.testing.assert_close(wg.cpu(), wc, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rg.cpu(), rc, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("cluster,loopback,slices", [(1, False, 2), (4, True, 2), (8, True, 1), (2, True, 2)])
def test_sgd_epoch_kernel_single_gpu(native_built, cluster, loopback, slices):
    """Whole-epoch launch (csrc/mf.cu::sgd_epoch_kernel): every rating exactly once per epoch, loss falls like the
    per-panel path's.  ``loopback`` runs the in-kernel rotation protocol (ready / consumed / pulled flags, comm
    warps, ping-pong buffers) as a ring of ONE rank whose peer slab is its own, so the driver's one-GPU box
    covers it; ``cluster`` > 1 exercises the cluster-of-CTAs workers and their DSMEM hand-over."""
    from harp_b200 import _native
    from harp_b200.models.mfsgd import MFSGDEngine
    r, c, v = _problem(nr=20000, nc=3000, nnz=600000, rank=64, seed=2)
    r, c, v = r.cuda(), c.cuda(), v.cuda()
    eng = MFSGDEngine(r, c, v, 20000, 3000, 64, 0.01, 0.02, seed=5, cluster=cluster, loopback=loopback, num_slices=slices,
                      fused=True)
    assert eng.fused and eng.cs == cluster
    ref = MFSGDEngine(r, c, v, 20000, 3000, 64, 0.01, 0.02, seed=5, fused=False, num_slices=slices)
    e0 = eng.rmse(r, c, v)
    _native.reset_launches()
    for it in range(8):
        assert int(eng.epoch().item()) == r.numel()
        assert int(ref.epoch().item()) == r.numel()
    eng.synchronize(); ref.synchronize()
    assert _native.launches() >= 8
    e1, f1 = eng.rmse(r, c, v), ref.rmse(r, c, v)
    assert math.isfinite(e1) and e1 < 0.5 * e0, (e0, e1)
    assert abs(e1 - f1) < 0.1 * f1, (e1, f1)            # same algorithm, different interleaving
    for per in eng.panels:
        for p in per:
            assert int(p.col_ticket.min()) == int(p.col_ticket.max()) == 8 * eng.B
    if loopback:
        words = eng.ring.state[:64].view(torch.int32).cpu().tolist()
        G = eng.B * eng.cs
        assert words[10] == 0                                         # no wait expired
        assert words[0] == 8 * G and words[2] == 8 * G                # done / pulled of slice 0
        assert words[6] == 8 and words[8] == 8                        # ready / consumed of slice 0


def test_sgd_epoch_deadline_and_collision_bound(native_built):
    from harp_b200.models.mfsgd import MFSGDEngine
    from harp_b200.ops import mf
    r, c, v = _problem(nr=200000, nc=20000, nnz=4_000_000, rank=128, seed=5)
    r, c, v = r.cuda(), c.cuda(), v.cuda()
    eng = MFSGDEngine(r, c, v, 200000, 20000, 128, 0.001, 0.01, seed=1, loopback=True)
    full = int(eng.epoch().item())
    part = int(eng.epoch(deadline_ms=0.02).item())
    eng.synchronize()
    assert full == 4_000_000 and 0 <= part < full
    # rows that two warps of one worker may touch at the same time (different column runs of one cell)
    frac = mf.row_collision_bound(eng.panels[0][0])
    print(f"row exposure bound of panel (0,0): {frac:.4f}")
    assert 0.0 <= frac <= 1.0, frac


@pytest.mark.parametrize("nr,nc,nnz", [(3000, 400, 60000), (20000, 20, 200000)])
def test_ccd_engine_device_matches_cpu(native_built, nr, nc, nnz):
    """csrc/mf.cu::ccd_sweep_kernel (warp per id, and CTA per id for the long lists of the second shape) against the
    fp64 CPU engine: same RMSE trajectory and factors after three outer iterations."""
    from harp_b200 import _native
    from harp_b200.models.mfccd import MFCCDEngine
    g = torch.Generator().manual_seed(1)
    key = torch.unique(torch.randint(0, nr * nc, (nnz,), generator=g))
    rows, cols = key // nc, key % nc
    vals = (torch.rand(key.numel(), generator=g) * 4 + 1).double()
    cpu = MFCCDEngine(rows, cols, vals, rows, cols, vals, nr, nc, 6, 0.05, seed=2)
    gpu = MFCCDEngine(rows.cuda(), cols.cuda(), vals.cuda(), rows.cuda(), cols.cuda(), vals.cuda(), nr, nc, 6, 0.05, seed=2)
    gpu.W_loc.copy_(cpu.W_loc)                  # the device generator draws a different initial W
    assert gpu.wide_c == (nc == 20) and not gpu.wide_r
    l0 = _native.launches()
    for it in range(3):
        a, b = cpu.iterate(), gpu.iterate()
        assert abs(a - b) < 2e-4 * a, (it, a, b)
    assert _native.launches() - l0 == 3 * 6 * 2            # one kernel pass per side and feature, nothing else
    torch.testing.assert_close(gpu.W_loc.cpu().double(), cpu.W_loc, rtol=5e-3, atol=5e-4)
    torch.testing.assert_close(gpu.H_loc.cpu().double(), cpu.H_loc, rtol=5e-3, atol=5e-4)
