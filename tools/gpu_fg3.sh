#!/bin/bash
# GPU pass for final gathering: whole GPU suite, fg stage timings of both lookup variants at two sizes, ncu launch list, smoke.   usage: tools/gpu_fg3.sh <tag>
TAG=${1:-r03d}
O=gpurun_out/$TAG
mkdir -p $O
timeout 100 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 40 python tools/bench_fg.py > $O/bench_fg.jsonl 2> $O/bench_fg.err; echo "rc=$?" >> $O/bench_fg.err
B2RT_FG_VARIANT=1 timeout 40 python tools/bench_fg.py > $O/bench_fg_coop.jsonl 2> $O/bench_fg_coop.err; echo "rc=$?" >> $O/bench_fg_coop.err
timeout 40 python tools/bench_fg.py 8000000 2000000 1920 1080 32 1 > $O/bench_fg_cfg3.jsonl 2> $O/bench_fg_cfg3.err; echo "rc=$?" >> $O/bench_fg_cfg3.err
B2RT_FG_VARIANT=1 timeout 40 python tools/bench_fg.py 8000000 2000000 1920 1080 32 1 > $O/bench_fg_cfg3_coop.jsonl 2> $O/bench_fg_cfg3_coop.err; echo "rc=$?" >> $O/bench_fg_cfg3_coop.err
B2RT_FG_VARIANT=1 timeout 60 ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/launches_fg.csv \
    python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/ncu_fg.log 2>&1
timeout 30 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.txt 2>&1
ls -la $O
