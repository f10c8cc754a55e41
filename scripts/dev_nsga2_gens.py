import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from evox_myc_b200 import nsga2
for pop in (4096, 16384, 32768):
    algo = nsga2.NSGA2(np.zeros(12), np.ones(12), 3, pop); prob = nsga2.DTLZ2(m=3)
    st = algo.setup(0); x, st = algo.init_ask(st); f, _ = prob.evaluate(None, x); st = algo.init_tell(st, f)
    def gen(st):
        x, st = algo.ask(st); f, _ = prob.evaluate(None, x); return algo.tell(st, f)
    for _ in range(10): st = gen(st)
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): st = gen(st)
    e1.record(); torch.cuda.synchronize()
    print(pop, "ms/gen %.3f" % (e0.elapsed_time(e1) / 50))
