"""One warm-up + one measured step of the SHARDED entry point on a one-rank communicator (for ncu: never profile
a multi-rank command; the per-rank kernels are the same, a rank of W then streams 1/W of these bytes)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import torch.distributed as dist
from evox_myc_b200 import sharded

n, m = (int(v) for v in (sys.argv[1] if len(sys.argv) > 1 else "262144x5").split("x"))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29533")
dist.init_process_group("gloo", rank=0, world_size=1)
x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
sel = sharded.ShardedSelector(n, m, torch.device("cuda", 0))
for _ in range(2):
    o = sel.nsga2_select(x, n // 2)
torch.cuda.synchronize()
print("ok", n, m, int(o[0]), sel.last_path())
sel.close(); dist.destroy_process_group()
