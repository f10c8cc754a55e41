# Multi-GPU run: sharded parity tests on all exchange paths + stage times per path.
# usage: bash scripts/gpu_r2_multi.sh <tag> <world>    (under gpurun --gpus <world>)
TAG=${1:-r2multi}
W=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi -L | head -8
SEL=${3:-"${W}-data_flow or ${W}-auto or full_size and ${W}]"}
timeout 1500 python -m pytest tests/test_sharded_gpu.py -x -q -k "$SEL" > $OUT/tests_w$W.log 2>&1; echo "tests rc=$?"; tail -4 $OUT/tests_w$W.log
for mode in flow grid; do
  EVOXB_PEEL_MODE=$mode timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W --steps 20 --warmup 3 > $OUT/bench_${mode}_w$W.json 2> $OUT/bench_${mode}_w$W.err; echo "bench $mode rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench_${mode}_w$W.json").read().strip().splitlines()[-1])
    print("$mode", "ms_per_step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), d["stages_ms"])
except Exception as ex:
    print("no bench line:", ex); print(open("$OUT/bench_${mode}_w$W.err").read()[-1500:])
PY
done
