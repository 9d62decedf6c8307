"""Kernel-level throughput of LDA-CGS (synthetic Zipf corpus)."""
import json, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from harp_b200.models.lda import LDAEngine

D = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000
V = int(sys.argv[2]) if len(sys.argv) > 2 else 100_000
K = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
Ldoc = int(sys.argv[4]) if len(sys.argv) > 4 else 100
ntok = D * Ldoc
g = torch.Generator(device="cuda").manual_seed(0)
docs = torch.arange(D, device="cuda").repeat_interleave(Ldoc)
words = (torch.rand(ntok, device="cuda", generator=g) ** 3 * V).long() % V        # Zipf-like
import itertools
Ss = [int(x) for x in (sys.argv[6].split(",") if len(sys.argv) > 6 else ["2"])]
for B, S in itertools.product([int(x) for x in (sys.argv[5].split(",") if len(sys.argv) > 5 else ["148"])], Ss):
    eng = LDAEngine(docs, words, D, V, K, 0.01, 0.01, blocks=B, seed=1, num_slices=S)
    for _ in range(2): eng.iterate()
    eng.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): eng.iterate()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print(json.dumps({"B": eng.B, "S": eng.S, "tokens": ntok, "K": K, "ms_per_iter": ms, "tokens_per_s": ntok / ms * 1e3,
                      "ll": eng.log_likelihood()}))
    del eng
