"""One launch each of the K21 / K22 kernels at a mid size, for `ncu --set full -k regex:tsqr_kernel|solve_kernel`."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from harp_b200.ops import linalg as KL

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
X = torch.randn(4_000_000, 64, device=dev, generator=g)
for _ in range(2):
    KL.tsqr_r(X)
torch.cuda.synchronize()
rng = np.random.default_rng(0)
users = torch.sort(torch.from_numpy(rng.integers(0, 200_000, 5_000_000)).to(dev)).values
items = torch.from_numpy((rng.zipf(1.3, 5_000_000) % 100_000).astype(np.int64)).to(dev)
ptr = torch.zeros(200_001, dtype=torch.int64, device=dev); ptr[1:] = torch.cumsum(torch.bincount(users, minlength=200_000), 0)
Y = torch.rand(100_000, 64, device=dev, generator=g) * 0.1
out = torch.empty(200_000, 64, device=dev)
side = KL.ALSSide(ptr, items, torch.ones(5_000_000, device=dev))
for _ in range(2):
    KL.als_solve(side, Y, out, 64, 0.1, 40.0)
torch.cuda.synchronize()
print("NCU_LINALG_OK")
