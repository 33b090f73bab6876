#!/bin/bash
# photon shoot kernel: ncu full capture with source counters (two launches of photon_trace_regen_kernel)
TAG=${1:-shoot}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 ncu --clock-control none --set full --import-source on -k regex:'photon_trace_regen' -s 2 -c 2 -f -o $O/shoot_full \
    python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_shoot.log 2>&1
ls -la $O
