#!/bin/bash
# The ncu evidence of one round, all on the bench command itself (1 GPU; never under torchrun):
#   1. the plain run first (must exit 0 without ncu)
#   2. launch list: gpu__time_duration of every launch of OUR kernels -> per-kernel share of the step
#   3. DRAM traffic of the 12 launches the dominant kernel makes in one timed step (one Farneback batch per step)
#   4. --set full of three level-0 launches of the dominant kernel (source-level pages imported)
# usage: scripts/ncu_pass.sh <tag>      outputs under gpurun_out/<tag>_*
set -u
TAG=${1:-r2}
OUT=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline --min-seconds 0"
mkdir -p $OUT
$CMD > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err || { echo "plain run failed"; tail -5 $OUT/${TAG}_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k "regex:^k_|_t<" -c 4000 --csv \
    --log-file $OUT/${TAG}_ncu_launches_bench.csv $CMD > $OUT/${TAG}_ncu1.log 2>&1
python scripts/ncu_launches.py $OUT/${TAG}_ncu_launches_bench.csv > $OUT/${TAG}_ncu_launches_bench.txt
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
    -k regex:k_fb_blur_solve_t --launch-skip 36 -c 12 --csv --log-file $OUT/${TAG}_blur_traffic.csv $CMD > $OUT/${TAG}_ncu2.log 2>&1
# launch order: 3 warm-up steps x 12 launches, then per step levels 3, 2, 1, 0 x 3 launches: skip to the level-0 launches
# of the first timed step (one batch of all ~116 ROI crops of the step)
ncu --set full --clock-control none --import-source on -k regex:k_fb_blur_solve_t --launch-skip 45 -c 3 \
    -f -o $OUT/${TAG}_blur_full $CMD > $OUT/${TAG}_ncu3.log 2>&1
python scripts/ncu_summary.py $OUT/${TAG}_blur_full.ncu-rep > $OUT/${TAG}_ncu_full_blur_solve_bench.txt
ls -la $OUT | grep $TAG
head -14 $OUT/${TAG}_ncu_launches_bench.txt
cat $OUT/${TAG}_ncu_full_blur_solve_bench.txt | cut -c1-260
