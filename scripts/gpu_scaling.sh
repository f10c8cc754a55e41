# 1/2/4/8-GPU sweep of bench.py (same launch lines the driver uses) + multi-GPU parity tests
set -x
OUT=gpurun_out/${1:-scale}
mkdir -p $OUT
NG=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -x -q -k "${TESTSEL:-multi_gpu}" 2>&1 | tail -4 | tee $OUT/tests.log
python bench.py --gpus 1 --steps 20 --warmup 3 --no-cpu-baseline > $OUT/bench_n1.json 2> $OUT/bench_n1.err
for N in 2 4 8; do
  if [ $N -le $NG ]; then
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 3 > $OUT/bench_n$N.json 2> $OUT/bench_n$N.err
  fi
done
python - <<'PY'
import json,glob,os
out=os.environ.get('OUTDIR') or sorted(glob.glob('gpurun_out/*/bench_n1.json'))[-1].rsplit('/',1)[0]
base=None
for n in (1,2,4,8):
    p=f'{out}/bench_n{n}.json'
    if not os.path.exists(p) or not os.path.getsize(p): continue
    d=json.loads(open(p).read().strip().splitlines()[-1])
    base=base or d['ms_per_step']
    print(n, 'ms %.3f'%d['ms_per_step'], 'speedup %.2f'%(base/d['ms_per_step']), 'e2e ms %.3f'%d['e2e']['ms_per_step'], d['stages_ms'])
PY
