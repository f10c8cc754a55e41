# Round profile: bench lines + ncu launch list + ncu full capture of the three heavy kernels.
# usage: bash scripts/gpu_profile.sh <tag>   (run under gpurun, 1 GPU)
set -x
TAG=${1:-r01b}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python bench.py --steps 30 --warmup 3 > $OUT/bench_c4.json 2> $OUT/bench_c4.err; tail -c 400 $OUT/bench_c4.err
python bench.py --steps 30 --warmup 3 --workload c2 --no-cpu-baseline > $OUT/bench_c2.json 2>> $OUT/bench_c4.err
python bench.py --steps 30 --warmup 3 --workload c5 --no-cpu-baseline > $OUT/bench_c5.json 2>> $OUT/bench_c4.err
python bench.py --steps 30 --warmup 3 --workload c3 --no-cpu-baseline > $OUT/bench_c3.json 2>> $OUT/bench_c4.err
python scripts/profile_step.py 65536x3 > $OUT/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_c2.csv python scripts/profile_step.py 65536x3 > $OUT/ncu_c2.log 2>&1
python scripts/profile_step.py 65536x3 > $OUT/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'peel_persistent|dominance_build|dominator_count|presort_coop' -s 4 -c 4 -o $OUT/prof_c2 -f python scripts/profile_step.py 65536x3 > $OUT/ncu_c2_full.log 2>&1
python scripts/profile_step.py 262144x5 > $OUT/plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 200 --csv --log-file $OUT/launches_c4.csv python scripts/profile_step.py 262144x5 > $OUT/ncu_c4.log 2>&1
ls -la $OUT
