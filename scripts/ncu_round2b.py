"""One launch of each of the smaller round-2 kernels for one ncu pass:
    ncu --set full --clock-control none --import-source on -k regex:"solve_kernel|ccd_sweep|combine_kernel|bc_stress|potrs" -c 6 ..."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
from harp_b200.ops import linalg as KL, mf, subgraph as SG, mds as KM

rng = np.random.default_rng(0)
users = torch.sort(torch.from_numpy(rng.integers(0, 200_000, 5_000_000)).to(dev)).values
items = torch.from_numpy((rng.zipf(1.3, 5_000_000) % 100_000).astype(np.int64)).to(dev)
ptr = torch.zeros(200_001, dtype=torch.int64, device=dev); ptr[1:] = torch.cumsum(torch.bincount(users, minlength=200_000), 0)
Y = torch.rand(100_000, 64, device=dev, generator=g) * 0.1
out = torch.empty(200_000, 64, device=dev)
side = KL.ALSSide(ptr, items, torch.ones(5_000_000, device=dev))
KL.als_solve(side, Y, out, 64, 0.1, 40.0); torch.cuda.synchronize()

# CCD sweep over rows (warp per id) of a 10 M-rating share
nr, nc, nnz = 480_000, 17_700, 10_000_000
rows = torch.sort((torch.rand(nnz, device=dev, generator=g) ** 2 * nr).long() % nr).values
cols = ((torch.rand(nnz, device=dev, generator=g) ** 2 * nc).long() % nc).to(torch.int32)
rptr = torch.zeros(nr + 1, dtype=torch.int32, device=dev); rptr[1:] = torch.cumsum(torch.bincount(rows, minlength=nr), 0).to(torch.int32)
res = torch.rand(nnz, device=dev, generator=g)
u = torch.rand(nr, device=dev, generator=g); v = torch.rand(nc, device=dev, generator=g)
mf.ccd_sweep(rptr, cols, res, v, u, 0.1, pd_other=torch.rand(nc, device=dev, generator=g) * 0.01, pd_self=u.clone(), dself=torch.empty(nr, device=dev))
torch.cuda.synchronize()

# subgraph DP step: size-6 table from 3 + 3 over 10 colours, 1 M vertices, degree 8
n, deg, k = 1_000_000, 8, 10
nbr_ptr = torch.arange(n + 1, device=dev, dtype=torch.int64) * deg
nbr_idx = torch.randint(0, n, (n * deg,), device=dev, generator=g).to(torch.int32)
sa, sp = SG.split_tables(k, 6, 3, dev)
act = torch.rand(n, len(SG.colour_sets(k, 3)), device=dev, generator=g)
pas = torch.rand(n, len(SG.colour_sets(k, 3)), device=dev, generator=g)
SG.combine(nbr_ptr, nbr_idx, act, pas, sa, sp); torch.cuda.synchronize()

# MDS pass: 20 000 points, all rows local
m = 20_000
P = torch.rand(m, 3, device=dev, generator=g)
D = torch.cdist(P, P)
KM.bc_stress(D, torch.arange(m, device=dev), torch.rand(m, 3, device=dev, generator=g).double(), 0.0); torch.cuda.synchronize()
print("NCU_ROUND2B_OK")
