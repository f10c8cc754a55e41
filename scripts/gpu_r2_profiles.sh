# Round-2 evidence: bench lines of every workload + ncu launch lists (with the source hash beside them) + full ncu
# captures of the build, peel and data-flow peel kernels.   usage: bash scripts/gpu_r2_profiles.sh <tag> (1 GPU)
TAG=${1:-r02}
VER=${2:-v2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
HASH=$(python -c "import bench; print(bench.library_source_hash())")
python bench.py --steps 30 --warmup 3 > $OUT/bench_c4.json 2> $OUT/bench.err; echo "bench c4 rc=$?"
for w in c2 c3 c5; do python bench.py --steps 30 --warmup 3 --workload $w --no-cpu-baseline --no-nsga2 > $OUT/bench_$w.json 2>> $OUT/bench.err; done
python bench.py --steps 20 --warmup 3 --workload 65536x2 --no-cpu-baseline --no-nsga2 > $OUT/bench_deep_65536x2.json 2>> $OUT/bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference_c4.json 2>> $OUT/bench.err
for shape in 262144x5 65536x3; do
  N=${shape%x*}; M=${shape#*x}; W=c4; [ $N = 65536 ] && W=c2
  python scripts/profile_step.py $shape > $OUT/plain_$W.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_${W}_N${N}_M${M}_${VER}.csv python scripts/profile_step.py $shape > $OUT/ncu_$W.log 2>&1
  echo "{\"src_hash\": \"$HASH\", \"command\": \"ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python scripts/profile_step.py $shape\"}" > $OUT/launches_${W}_N${N}_M${M}_${VER}.csv.meta.json
done
EVOXB_PEEL_MODE=grid python scripts/profile_step_sharded.py 262144x5 > $OUT/plain_c4_sharded.log 2>&1 && EVOXB_PEEL_MODE=grid ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_c4sharded_N262144_M5_w1_${VER}.csv python scripts/profile_step_sharded.py 262144x5 > $OUT/ncu_c4_sharded.log 2>&1
echo "{\"src_hash\": \"$HASH\", \"command\": \"EVOXB_PEEL_MODE=grid ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none python scripts/profile_step_sharded.py 262144x5\"}" > $OUT/launches_c4sharded_N262144_M5_w1_${VER}.csv.meta.json
# full captures: gpurun brings back at most 64 MiB, a report with sources is 10 MB per kernel
python scripts/profile_step.py 262144x5 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'dominance_build_lvl|peel_persistent|level_fixup_find' -s 3 -c 3 -o $OUT/prof_c4 -f python scripts/profile_step.py 262144x5 > $OUT/ncu_c4_full.log 2>&1
if [ "${FULL:-0}" = "1" ]; then
python scripts/profile_nsga3.py > /dev/null 2>&1 && ncu --set full --clock-control none -k regex:'n3_associate|n3_niche' -s 2 -c 2 -o $OUT/prof_c5n3 -f python scripts/profile_nsga3.py > $OUT/ncu_c5n3_full.log 2>&1
python scripts/profile_step.py 65536x3 > /dev/null 2>&1 && ncu --set full --clock-control none -k regex:'peel_flow|presort_coop' -s 2 -c 2 -o $OUT/prof_c2 -f python scripts/profile_step.py 65536x3 > $OUT/ncu_c2_full.log 2>&1
fi
du -sh $OUT
ls -la $OUT; tail -c 300 $OUT/bench.err
