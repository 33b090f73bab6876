#!/bin/bash
# ncu --set full capture of the final-gather kernel on a small frame (the kernel runs ~20 ms there, so the ~40 replays fit)
TAG=${1:-r03g}
O=gpurun_out/$TAG
mkdir -p $O
timeout 70 ncu --clock-control none --set full --import-source on -k regex:final_gather_kernel -c 1 -f -o $O/fg_full \
    python tools/bench_fg.py 300000 30000 320 180 32 1 > $O/ncu_fg_full.log 2>&1
ls -la $O
