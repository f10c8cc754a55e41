# Last check of a kernel change: level / ranking parity tests, c4 and c5 bench lines, and the three ncu launch lists with
# the source hash (what bench.py reads roofline.traffic from).   usage: bash scripts/gpu_r2_oneshot.sh <tag> <ver>
TAG=${1:-r02e}
VER=${2:-v5}
OUT=gpurun_out/$TAG
mkdir -p $OUT
HASH=$(python -c "import bench; print(bench.library_source_hash())")
timeout 300 python -m pytest tests/test_levels_gpu.py tests/test_nds_gpu.py -x -q 2>&1 | tail -3 | tee $OUT/tests.txt
for w in c4 c5; do python bench.py --steps 20 --warmup 3 --workload $w --no-cpu-baseline --no-nsga2 > $OUT/bench_$w.json 2>> $OUT/bench.err; python - <<PY
import json
d=json.loads(open("$OUT/bench_$w.json").read().strip().splitlines()[-1])
print("$w", round(d["ms_per_step"],3), d.get("stages_ms"))
PY
done
for shape in 262144x5 65536x3; do
  N=${shape%x*}; M=${shape#*x}; W=c4; [ $N = 65536 ] && W=c2
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_${W}_N${N}_M${M}_${VER}.csv python scripts/profile_step.py $shape > $OUT/ncu_$W.log 2>&1
  echo "{\"src_hash\": \"$HASH\", \"command\": \"ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python scripts/profile_step.py $shape\"}" > $OUT/launches_${W}_N${N}_M${M}_${VER}.csv.meta.json
done
EVOXB_PEEL_MODE=grid ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_c4sharded_N262144_M5_w1_${VER}.csv python scripts/profile_step_sharded.py 262144x5 > $OUT/ncu_c4_sharded.log 2>&1
echo "{\"src_hash\": \"$HASH\", \"command\": \"EVOXB_PEEL_MODE=grid ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python scripts/profile_step_sharded.py 262144x5\"}" > $OUT/launches_c4sharded_N262144_M5_w1_${VER}.csv.meta.json
ls $OUT | wc -l
