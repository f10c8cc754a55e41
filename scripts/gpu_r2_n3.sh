# NSGA-III tests + generation-loop timings (NSGA-II per peel mode, NSGA-III stages)
OUT=gpurun_out/${1:-r2n3}; mkdir -p $OUT
timeout 900 python -m pytest tests/test_nsga3_gpu.py -x -q 2>&1 | tail -3
for mode in auto grid flow; do
  if [ $mode = auto ]; then unset EVOXB_PEEL_MODE; else export EVOXB_PEEL_MODE=$mode; fi
  echo "nsga2 gens, peel mode $mode"; timeout 300 python scripts/dev_nsga2_gens.py 2>&1 | tail -3
done
unset EVOXB_PEEL_MODE
python - <<PY
import json, types, torch, bench
args = types.SimpleNamespace(steps=20)
print(json.dumps(bench.bench_nsga3(args, torch.device("cuda", 0))))
PY
