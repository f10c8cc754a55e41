# The driver's multi-GPU bench line (auto mode) on W GPUs.  usage: bash scripts/gpu_r2_bench_multi.sh <tag> <W>
TAG=${1:-r2bm}; W=${2:-2}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $W --steps 20 --warmup 3 > $OUT/bench_w$W.json 2> $OUT/bench_w$W.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/bench_w$W.json").read().strip().splitlines()[-1])
    print("ms_per_step", round(d["ms_per_step"], 3), "best", round(d["ms_per_step_best"], 3), "e2e", round(d["e2e"]["ms_per_step"], 3), d["stages_ms"])
    print("exchange", d["exchange"]); print("parity", d["parity"]); print("roofline kernel", d["roofline"]["kernel"], d["roofline"]["frac"])
except Exception as ex:
    print("no bench line:", ex); print(open("$OUT/bench_w$W.err").read()[-2500:])
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus $W --steps 2 --warmup 1 2>/dev/null | python -c "
import sys, json
for ln in sys.stdin:
    try: d = json.loads(ln); print('reference arm', d['value'], d['cpu_baseline'])
    except Exception: pass
"
