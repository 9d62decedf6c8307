"""One launch of every round-2 hot kernel at a representative size, for ONE ncu pass:

    ncu --set full --clock-control none --import-source on -k regex:"epoch_kernel|assign_tc_pair|tsqr_kernel|solve_kernel" \
        -c 7 -o gpurun_out/round2 python scripts/ncu_round2.py

The epoch kernels run with one CTA per worker (ncu cannot replay cooperative launches of thread-block clusters:
LaunchFailed), on one rank's share of an 8-rank job with the loop-back ring, so the in-kernel rotation is active."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)

from harp_b200.models.kmeans import KMeansEngine
X = torch.rand(2_000_000, 128, device=dev, generator=g)
eng = KMeansEngine(X, 10_000, precision="fp16x3")
eng.set_centroids(torch.rand(10_000, 128, device=dev, generator=g))
eng.step(); torch.cuda.synchronize()
del eng, X

from harp_b200.models.mfsgd import MFSGDEngine
N = 8
rows, cols, nnz = 480_000 // N, 17_700, 100_000_000 // N
r = (torch.rand(nnz, device=dev, generator=g) ** 2 * rows).long() % rows
c = (torch.rand(nnz, device=dev, generator=g) ** 2 * cols).long() % cols
v = torch.rand(nnz, device=dev, generator=g) * 4 + 1
sgd = MFSGDEngine(r, c, v, rows, cols, 128, 0.001, 0.01, seed=3, virtual_world=N, cluster=1)
sgd.epoch(); sgd.synchronize()
del sgd, r, c, v

from harp_b200.models.lda import LDAEngine
D, V, K, L = 1_000_000 // N, 100_000, 1000, 100
docs = torch.arange(D, device=dev).repeat_interleave(L)
words = (torch.rand(D * L, device=dev, generator=g) ** 3 * V).long() % V
lda = LDAEngine(docs, words, D, V, K, 0.01, 0.01, seed=1, cluster=1, virtual_world=N)
lda.iterate(); lda.synchronize()
del lda
tl = LDAEngine(docs[: 20_000 * L], words[: 20_000 * L], 20_000, V, 10_000, 0.01, 0.01, seed=1, cluster=1, sampler="tokenlist", loopback=True)
tl.iterate(); tl.synchronize()
del tl, docs, words

from harp_b200.ops import linalg as KL
Xq = torch.randn(4_000_000, 64, device=dev, generator=g)
KL.tsqr_r(Xq); torch.cuda.synchronize()
rng = np.random.default_rng(0)
users = torch.sort(torch.from_numpy(rng.integers(0, 200_000, 5_000_000)).to(dev)).values
items = torch.from_numpy((rng.zipf(1.3, 5_000_000) % 100_000).astype(np.int64)).to(dev)
ptr = torch.zeros(200_001, dtype=torch.int64, device=dev); ptr[1:] = torch.cumsum(torch.bincount(users, minlength=200_000), 0)
Y = torch.rand(100_000, 64, device=dev, generator=g) * 0.1
out = torch.empty(200_000, 64, device=dev)
side = KL.ALSSide(ptr, items, torch.ones(5_000_000, device=dev))
KL.als_solve(side, Y, out, 64, 0.1, 40.0); torch.cuda.synchronize()
print("NCU_ROUND2_OK")
