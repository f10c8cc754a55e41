# ncu capture of the data-flow peel kernel at c2 (one launch, full set)
TAG=${1:-r2ncu}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export EVOXB_PEEL_MODE=flow
python scripts/profile_step.py 65536x3 > $OUT/plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'peel_flow' -s 1 -c 1 -o $OUT/prof_flow_c2 -f python scripts/profile_step.py 65536x3 > $OUT/ncu.log 2>&1
tail -3 $OUT/ncu.log
ls -la $OUT
