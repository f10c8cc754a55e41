# Single-GPU check: full GPU tests, stage times over shapes, default bench line.
TAG=${1:-r2check}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 2000 python -m pytest tests -m gpu -x -q > $OUT/tests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/tests.log
timeout 600 python scripts/dev_stage_times.py 32768x3 65536x3 65536x2 65536x10 262144x5 > $OUT/stages.log 2>&1; grep -v "^---" $OUT/stages.log | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln)
    except Exception: print(ln.strip()); continue
    print(d['n'], d['m'], 'fronts', d.get('fronts'), 'total', d['total_ms'], d['stages_ms'])
"
timeout 900 python bench.py > $OUT/bench_c4.json 2> $OUT/bench_c4.err; echo "bench rc=$?"; tail -c 600 $OUT/bench_c4.err
python - <<PY
import json
d = json.loads(open("$OUT/bench_c4.json").read().strip().splitlines()[-1])
print({k: d[k] for k in ("value", "ms_per_step", "ms_per_step_best", "ms_per_step_median", "gpu_launches", "stages_ms")})
print("e2e", d["e2e"]); print("roofline", d["roofline"]); print("nsga2", d["nsga2"]); print("cpu", d["cpu_baseline"])
PY
