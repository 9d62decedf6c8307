"""Kernel-level timing of the Gram kernel vs torch (cuBLAS TF32)."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from harp_b200.ops.gram import gram

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
X = torch.randn(n, d, device="cuda")
G = torch.empty(d, d, device="cuda")
for _ in range(3): gram(X, out=G)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(); gram(X, out=G); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[2]
torch.backends.cuda.matmul.allow_tf32 = True
for _ in range(2): R = X.t() @ X
torch.cuda.synchronize()
e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record(); R = X.t() @ X; e1.record(); torch.cuda.synchronize()
out = {"n": n, "d": d, "ms": ms, "tflops_full": 2.0 * n * d * d / ms / 1e9, "torch_tf32_ms": e0.elapsed_time(e1),
       "max_rel_err_vs_torch": float((G - R).abs().max() / R.abs().max())}
print(json.dumps(out))
