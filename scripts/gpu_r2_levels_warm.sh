# Per-kernel times of one c4 step with warm caches (ncu --cache-control none), level build.  usage: <tag>  (1 GPU)
TAG=${1:-r02lvw}
OUT=gpurun_out/$TAG
mkdir -p $OUT
for shape in 262144x5 65536x10; do
  ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -c 200 --csv --log-file $OUT/launches_warm_$shape.csv python scripts/profile_step.py $shape > $OUT/ncu_$shape.log 2>&1
done
ls $OUT
