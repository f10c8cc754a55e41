# Level build: parity tests, bench lines of c4 / c5 / c2 with the level build and with the float build, and the
# ncu launch list of one step with levels.   usage: bash scripts/gpu_r2_levels.sh <tag>   (1 GPU)
TAG=${1:-r02lv}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 1500 python -m pytest tests/test_levels_gpu.py tests/test_nds_gpu.py -x -q 2>&1 | tail -15 | tee $OUT/tests.txt
for mode in levels exact; do
  for w in c4 c5 c2; do
    EVOXB_BUILD_MODE=$mode python bench.py --steps 20 --warmup 3 --workload $w --no-cpu-baseline --no-nsga2 > $OUT/bench_${w}_$mode.json 2>> $OUT/bench.err
    python - <<PY
import json
d=json.loads(open("$OUT/bench_${w}_$mode.json").read().strip().splitlines()[-1])
print("$w $mode", round(d["ms_per_step"],3), d.get("stages_ms"))
PY
  done
done
for shape in 262144x5 65536x10; do
  EVOXB_BUILD_MODE=levels ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_levels_$shape.csv python scripts/profile_step.py $shape > $OUT/ncu_$shape.log 2>&1
done
tail -c 400 $OUT/bench.err
