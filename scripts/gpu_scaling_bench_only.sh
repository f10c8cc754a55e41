# Scaling sweep, bench only (no tests): one bench line per GPU count under gpurun_out/<tag>/.
# usage (inside gpurun --gpus 8): bash scripts/gpu_scaling_bench_only.sh [tag] [gpu counts...]
TAG=${1:-scale}; shift || true
NS=${@:-1 2 4 8}
OUT=gpurun_out/$TAG; mkdir -p $OUT
for N in $NS; do
  if [ "$N" = 1 ]; then
    python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline --no-nsga2 > $OUT/bench_n1.json 2> $OUT/bench_n1.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
  fi
done
TAG=$TAG python - <<'PY'
import json, os
tag = os.environ['TAG']
for n in (1, 2, 4, 8):
    p = f'gpurun_out/{tag}/bench_n{n}.json'
    if not os.path.exists(p) or not os.path.getsize(p):
        continue
    d = json.loads(open(p).read().strip().splitlines()[-1])
    print(n, 'ms %.3f' % d['ms_per_step'], 'e2e ms %.3f' % d['e2e']['ms_per_step'], d['stages_ms'])
PY
