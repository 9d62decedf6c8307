import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from harp_b200.ops.gram import gram
torch.manual_seed(0)
for n in (4096, 200000, 2000000):
    d = 1024
    X = torch.randn(n, d, device="cuda")
    G = gram(X); torch.cuda.synchronize()
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = (X.t() @ X).double() if n > 500000 else X.double().t() @ X.double()
    err = (G.double() - ref).abs()
    tiles = err.view(4, 256, 4, 256).amax(dim=(1, 3)) / ref.abs().max()
    print("n", n, "rel err per macro tile:\n", tiles.cpu().numpy().round(5))
    print(" diag ratio min/max", float((G.diag().double() / ref.diag()).min()), float((G.diag().double() / ref.diag()).max()))
