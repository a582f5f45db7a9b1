#!/bin/bash
# Round profile capture, run on the GPU box:  gpurun --timeout 900 -- 'bash tools/ncu_capture.sh r01_v4'
# 1) bench without ncu must exit 0; 2) launch list of the same command; 3) --set full of the decode and extend
# kernels (2 launches each).  Everything lands in gpurun_out/; summaries are made here by profiles/ncu_summary.py.
set -u
TAG=${1:-cap}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --headline-only --no-cpu-baseline --steps 4 --warmup 3"
timeout 300 $CMD > $OUT/${TAG}_bench_plain.json 2> $OUT/${TAG}_bench_plain.err || { echo "plain bench failed"; tail -5 $OUT/${TAG}_bench_plain.err; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 6 -c 2 \
    -f -o $OUT/${TAG}_decode_full $CMD > $OUT/${TAG}_ncu_decode.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extend_kernel -s 4 -c 2 \
    -f -o $OUT/${TAG}_extend_full $CMD > $OUT/${TAG}_ncu_extend.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:extend_kernel -s 2 -c 1 \
    -f -o $OUT/${TAG}_extend_cfg4_full python tools/prof_cfg4.py > $OUT/${TAG}_ncu_extend_cfg4.log 2>&1
ls -la $OUT | tail -12
