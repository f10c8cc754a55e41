# Multi-GPU bench line with a forced peel mode.  usage: bash scripts/gpu_r2_bench_mode.sh <tag> <W> <mode>
TAG=${1:-r2bm}; W=${2:-2}; MODE=${3:-flow}; OUT=gpurun_out/$TAG; mkdir -p $OUT
EVOXB_PEEL_MODE=$MODE timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $W --steps 20 --warmup 3 > $OUT/bench_${MODE}_w$W.json 2> $OUT/bench_${MODE}_w$W.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench_${MODE}_w$W.json").read().strip().splitlines()[-1])
    print("$MODE", "ms_per_step", round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), d["stages_ms"], d["exchange"]["path"], d["parity"])
except Exception as ex:
    print("no bench line:", ex); print(open("$OUT/bench_${MODE}_w$W.err").read()[-2000:])
PY
