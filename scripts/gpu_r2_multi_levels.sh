# Multi-GPU check of the level build: sharded parity tests (auto, levels, full size) + the driver's bench line.
# usage: bash scripts/gpu_r2_multi_levels.sh <tag> <world>    (under gpurun --gpus <world>)
TAG=${1:-r2ml}
W=${2:-2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1200 python -m pytest tests/test_sharded_gpu.py -x -q -k "${W}-levels or ${W}-auto or full_size and ${W}]" > $OUT/tests_w$W.log 2>&1; echo "tests rc=$?"; tail -4 $OUT/tests_w$W.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W --steps 20 --warmup 3 > $OUT/scaling_n$W.json 2> $OUT/bench_w$W.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/scaling_n$W.json").read().strip().splitlines()[-1])
    print("ms_per_step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), d["stages_ms"], "parity", d.get("parity"))
except Exception as ex:
    print("no bench line:", ex); print(open("$OUT/bench_w$W.err").read()[-1500:])
PY
