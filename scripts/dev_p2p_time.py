import os, sys, torch, numpy as np
sys.path.insert(0, os.getcwd())
import torch.distributed as dist
rank = int(os.environ["RANK"]); torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
from evox_myc_b200 import sharded
n, m = 262144, 5
x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
sel = sharded.ShardedSelector(n, m)
for i in range(3):
    r = sel.non_dominated_sort(x); torch.cuda.synchronize()
print(sel.status())
dist.destroy_process_group()
