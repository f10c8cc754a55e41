set -x
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; tail -c 600 gpurun_out/bench_c4.err
python bench.py --steps 10 --warmup 3 --workload c2 --no-cpu-baseline > gpurun_out/bench_c2.json 2>> gpurun_out/bench_c4.err
python scripts/profile_step.py 65536x3 > gpurun_out/plain_c2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python scripts/profile_step.py 65536x3 > gpurun_out/ncu_c2.log 2>&1
python scripts/profile_step.py 65536x3 > gpurun_out/plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'peel_persistent|dominance_build|dominator_count' -s 3 -c 3 -o gpurun_out/prof_c2 -f python scripts/profile_step.py 65536x3 > gpurun_out/ncu_c2_full.log 2>&1
python scripts/profile_step.py 262144x5 > gpurun_out/plain_c4.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'peel_persistent|dominance_build|dominator_count' -s 3 -c 3 --csv --log-file gpurun_out/dram_c4.csv python scripts/profile_step.py 262144x5 > gpurun_out/ncu_c4.log 2>&1
ls -la gpurun_out
