# Peel kernel choice: stage times of non_dominated_sort over shapes, per peel mode.
TAG=${1:-r2sweep}
OUT=gpurun_out/$TAG
mkdir -p $OUT
SHAPES="16384x3 32768x3 65536x3 131072x3 131072x2 32768x4 65536x4 131072x4 32768x5 65536x5 65536x10 262144x5"
for mode in flow grid auto; do
  if [ $mode = auto ]; then unset EVOXB_PEEL_MODE; else export EVOXB_PEEL_MODE=$mode; fi
  timeout 600 python scripts/dev_stage_times.py $SHAPES 2>&1 | grep -v "^---" | head -12 | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln)
    except Exception: print(ln.strip()); continue
    print('$mode', d['n'], d['m'], 'fronts', d['fronts'], 'total', d['total_ms'], 'peel', d['stages_ms'].get('peel'))
" | tee -a $OUT/sweep.log
done
unset EVOXB_PEEL_MODE
timeout 900 python -m pytest tests/test_nds_gpu.py tests/test_random_differential_gpu.py tests/test_select_gpu.py -x -q 2>&1 | tail -3
