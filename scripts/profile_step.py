"""One warm-up step + one measured step of the hot path, for ncu (never a bench number)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import evox_myc_b200 as e

n, m = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "65536x3").split("x"))
x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
for _ in range(2):
    o = e.nsga2_select(x, n // 2)
torch.cuda.synchronize()
print("ok", n, m, int(o[0]))
