#!/bin/bash
# One GPU-box pass for the final-gathering work: its parity tests first, then the whole GPU suite, the fg stage timings,
# the bench (both arms) and an ncu launch list of the fg frame.   usage: tools/gpu_fg.sh <tag>
TAG=${1:-r03a}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/smi.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_zfinalgather.py -m gpu -q > $O/pytest_fg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_fg.log
timeout 400 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 120 python tools/bench_fg.py > $O/bench_fg.jsonl 2> $O/bench_fg.err; echo "rc=$?" >> $O/bench_fg.err
timeout 60 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.txt 2>&1
timeout 300 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 200 ncu --clock-control none --metrics gpu__time_duration.sum -c 600 --csv --log-file $O/launches_fg.csv \
    python tools/bench_fg.py 200000 20000 480 270 32 1 > $O/ncu_fg.log 2>&1
timeout 120 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
ls -la $O
tail -5 $O/pytest_fg.log $O/pytest_gpu.log
