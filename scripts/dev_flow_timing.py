"""Development aid: per-phase clock totals of every CTA of peel_flow_kernel (needs a library built with
-DEVOXB_FLOW_TIMING=1, see scripts/build_timing_lib.sh; run with EVOXB_LIB_PATH pointing at it)."""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import evox_myc_b200 as e
from evox_myc_b200 import _lib

L = _lib.lib()
names = ["setup", "flag", "wait flags", "read words", "compaction", "row loads", "level 1", "merge+column"]
cfgs = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(65536, 3)]
for n, m in cfgs:
    x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
    for _ in range(3):
        r = e.non_dominated_sort(x)
    out = (ctypes.c_longlong * (320 * 8 + 8))()
    L.evoxb_flow_timing_read(out)
    grid = int(out[320 * 8])
    a = np.array(out[: grid * 8], dtype=np.float64).reshape(grid, 8)
    fronts = int(r.max().item()) + 1
    work = a.sum(1) - a[:, 2]
    slow = int(work.argmax())
    print(json.dumps({"n": n, "m": m, "fronts": fronts, "grid": grid, "clocks_per_front": round(a.sum(1).mean() / (fronts + 1)),
                      "mean": {k: round(v / (fronts + 1)) for k, v in zip(names, a.mean(0))},
                      "max": {k: round(v / (fronts + 1)) for k, v in zip(names, a.max(0))},
                      "slowest_cta": slow, "slowest": {k: round(v / (fronts + 1)) for k, v in zip(names, a[slow])},
                      "work_by_cta_decile": [round(v / (fronts + 1)) for v in np.percentile(work, [0, 10, 50, 90, 100])]}))
