# Round-2 development run: parity tests of the ranking path, then stage times with the data-flow peel.
# usage: bash scripts/gpu_r2_flow.sh <tag>   (under gpurun, 1 GPU)
TAG=${1:-r2a}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export EVOXB_PEEL_MODE=flow
timeout 900 python -m pytest tests/test_nds_gpu.py tests/test_random_differential_gpu.py tests/test_select_gpu.py -x -q > $OUT/tests_flow.log 2>&1; echo "tests(flow) rc=$?"; tail -5 $OUT/tests_flow.log
timeout 300 python scripts/dev_stage_times.py 32768x3 65536x3 65536x2 65536x10 131072x4 262144x5 2>&1 | grep -v "^---" | head -6 > $OUT/stages_flow.log; cat $OUT/stages_flow.log
EVOXB_LIB_PATH=$PWD/build/libevoxb_timing.so timeout 300 python scripts/dev_flow_timing.py 32768x3 65536x3 65536x2 > $OUT/timing_flow.log 2>&1; cat $OUT/timing_flow.log
