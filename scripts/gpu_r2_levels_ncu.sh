# Full ncu capture of the level build kernel at c4 (and c5).   usage: bash scripts/gpu_r2_levels_ncu.sh <tag>  (1 GPU)
TAG=${1:-r02lvn}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python scripts/profile_step.py 262144x5 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'dominance_build_lvl|level_scan|level_fixup' -s 3 -c 3 -o $OUT/prof_c4_lvl -f python scripts/profile_step.py 262144x5 > $OUT/ncu_c4_lvl.log 2>&1
ls -la $OUT
