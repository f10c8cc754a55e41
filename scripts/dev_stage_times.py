"""Development aid: per-stage CUDA-event times of non_dominated_sort at a few sizes."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import evox_myc_b200 as e
from evox_myc_b200 import _lib

cfgs = [(16384, 3), (65536, 3), (65536, 2), (65536, 10), (262144, 5)]
if len(sys.argv) > 1:
    cfgs = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
_lib.profile_enable(True)
for n, m in cfgs:
    x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
    for it in range(3):
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record(); r = e.non_dominated_sort(x); t1.record(); torch.cuda.synchronize()
        st = _lib.profile_read()
    tot = t0.elapsed_time(t1)
    fronts = int(r.max().item()) + 1
    peel_gbs = n * n / 8 / (st["peel"] * 1e-3) / 1e9
    cnt_gbs = n * n / 8 / (st["dominator_count"] * 1e-3) / 1e9
    print(json.dumps({"n": n, "m": m, "fronts": fronts, "total_ms": round(tot, 3),
                      "stages_ms": {k: round(v, 3) for k, v in st.items()},
                      "pairs_per_s": n * n / (tot * 1e-3), "peel_GBs": round(peel_gbs, 1), "count_GBs": round(cnt_gbs, 1),
                      "build_pairs_per_s": n * n / (st["dominance_build"] * 1e-3)}))

print("--- nsga2_select (NDS + crowding + selection sort) ---")
for n, m in cfgs:
    x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
    for it in range(3):
        torch.cuda.synchronize()
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        t0.record(); o = e.nsga2_select(x, n // 2); t1.record(); torch.cuda.synchronize()
        st = _lib.profile_read()
    print(json.dumps({"n": n, "m": m, "total_ms": round(t0.elapsed_time(t1), 3), "stages_ms": {k: round(v, 3) for k, v in st.items()}}))
