"""Asks the CUDA driver whether NVLS multicast objects can be created on this box (one line per rank)."""
import ctypes, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from harp_b200 import _native
lr = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(lr)
lib = _native.lib()
lib.hb_nvls_probe.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.c_char_p, ctypes.c_int]
buf = ctypes.create_string_buffer(1024)
for w in sorted({world, torch.cuda.device_count()}):
    rc = lib.hb_nvls_probe(w, 64 << 20, buf, 1024)
    print(f"rank {lr}: world {w}: rc={rc} {buf.value.decode()}", flush=True)
