"""Development aid: NSGA-II generation loop (DTLZ2, pop 16384): ms per generation and the stage times of tell."""
import os, sys, json
root = sys.argv[1] if len(sys.argv) > 1 else os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import numpy as np, torch
from evox_myc_b200 import nsga2, _lib
pop = 16384
algo = nsga2.NSGA2(np.zeros(12), np.ones(12), 3, pop); prob = nsga2.DTLZ2(d=12, m=3)
st = algo.setup(0); x, st = algo.init_ask(st); f, _ = prob.evaluate(None, x); st = algo.init_tell(st, f)
def gen(st):
    x, st = algo.ask(st); f, _ = prob.evaluate(None, x); return algo.tell(st, f)
for _ in range(5): st = gen(st)
torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(30): st = gen(st)
e1.record(); torch.cuda.synchronize()
_lib.profile_enable(True)
acc = {}
for _ in range(5):
    st = gen(st)
    for k, v in _lib.profile_read().items(): acc[k] = acc.get(k, 0) + v / 5
print(root, "ms/gen %.3f" % (e0.elapsed_time(e1) / 30), {k: round(v, 3) for k, v in acc.items()})
