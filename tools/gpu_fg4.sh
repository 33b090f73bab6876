#!/bin/bash
# Last confirmation pass: final-gathering parity tests + fg stage timings (default lookup variant).   usage: tools/gpu_fg4.sh <tag>
TAG=${1:-r03f}
O=gpurun_out/$TAG
mkdir -p $O
timeout 60 python -m pytest tests/test_gpu_zfinalgather.py -m gpu -q > $O/pytest_fg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_fg.log
timeout 30 python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/bench_fg.jsonl 2> $O/bench_fg.err; echo "rc=$?" >> $O/bench_fg.err
timeout 30 python tools/bench_fg.py 8000000 2000000 1920 1080 32 1 > $O/bench_fg_cfg3.jsonl 2> $O/bench_fg_cfg3.err; echo "rc=$?" >> $O/bench_fg_cfg3.err
ls -la $O
