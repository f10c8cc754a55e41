"""Small end-to-end run for compute-sanitizer (memcheck): every kernel of the path at odd sizes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
import evox_myc_b200 as e
from evox_myc_b200 import nsga2
import helpers as h
from oracle import c_oracle

for n, m in [(9001, 4), (4097, 2), (130, 1), (5000, 18)]:
    x = h.nasty(n, m, n)
    order, rank, crowd = e.nsga2_select(x, n // 2, return_details=True)
    wo, wr, wc = c_oracle.nsga2_select(x, n // 2)
    assert np.array_equal(rank, wr) and np.array_equal(order, wo)
    d = e.crowding_distance(x, rank == 1)
    p, f = e.non_dominate(np.zeros((n, 3), np.float32), x, n // 3)
algo = nsga2.NSGA2(np.zeros(12), np.ones(12), 3, 333)
st = nsga2.run(algo, nsga2.DTLZ2(m=3), 3, key=1)
torch.cuda.synchronize()
print("sanitize run ok")
