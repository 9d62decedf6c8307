import torch, sys
from pathlib import Path; sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from harp_b200.ops import mds as KM
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
m = 20000
P = torch.rand(m, 3, device=dev, generator=g)
D = torch.cdist(P, P)
ids = torch.arange(m, device=dev)
X = torch.rand(m, 3, device=dev, generator=g).double()
for _ in range(3): KM.bc_stress(D, ids, X, 0.0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): KM.bc_stress(D, ids, X, 0.0)
e1.record(); torch.cuda.synchronize()
print("mds pass ms", e0.elapsed_time(e1) / 20, "GB/s", m * m * 4 / (e0.elapsed_time(e1) / 20) / 1e6)
