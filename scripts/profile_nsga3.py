"""One NSGA-III selection at BASELINE config 5 (for ncu; never a bench number)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import helpers as h
from evox_myc_b200 import nsga3
n, m, pop = 65536, 10, 32768
x = torch.from_numpy(h.lsmop1(n, m, seed=1)).cuda()
ref = torch.from_numpy(nsga3.UniformSampling(pop, m)()[0]).cuda()
for _ in range(2):
    idx, key = nsga3.nsga3_select(x, pop, ref, (0, 1))
torch.cuda.synchronize()
print("ok", int(idx[0]))
