#!/bin/bash
# Second GPU pass for final gathering (after the hierarchical radiance-photon lookup): parity tests, the whole GPU suite,
# fg stage timings at two sizes, an ncu launch list and one full capture of the gather kernel.   usage: tools/gpu_fg2.sh <tag>
TAG=${1:-r03b}
O=gpurun_out/$TAG
mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_zfinalgather.py -m gpu -q > $O/pytest_fg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_fg.log
timeout 300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
S0=$(date +%s)
timeout 100 python tools/bench_fg.py > $O/bench_fg.jsonl 2> $O/bench_fg.err; echo "rc=$?" >> $O/bench_fg.err
S1=$(date +%s)
if [ $((S1 - S0)) -lt 30 ]; then
  timeout 100 python tools/bench_fg.py 8000000 2000000 1920 1080 32 2 > $O/bench_fg_cfg3.jsonl 2> $O/bench_fg_cfg3.err; echo "rc=$?" >> $O/bench_fg_cfg3.err
  timeout 120 ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/launches_fg.csv \
      python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/ncu_fg.log 2>&1
  timeout 150 ncu --clock-control none --set full --import-source on -k regex:final_gather_kernel -c 1 -f -o $O/fg_full \
      python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/ncu_fg_full.log 2>&1
fi
ls -la $O
