#!/bin/bash
# traversal variants A/B (+ parity under each)
TAG=${1:-t}
O=gpurun_out/$TAG
mkdir -p $O
{
python tools/ab_trace.py
for v in 1 2; do for rf in 0 12 20; do B2RT_TRACE_VARIANT=$v B2RT_TRACE_REFILL=$rf python tools/ab_trace.py; done; done
} > $O/ab_trace.log 2>&1
B2RT_TRACE_VARIANT=2 B2RT_TRACE_REFILL=12 timeout 600 python -m pytest tests -m gpu -x -q -k "parity" > $O/pytest_v2.log 2>&1; echo "rc=$?" >> $O/pytest_v2.log
NCU="ncu --clock-control none"
B2RT_TRACE_VARIANT=2 B2RT_TRACE_REFILL=12 timeout 900 $NCU --set full --import-source on -k regex:trace_spec -s 8 -c 2 -f -o $O/trace_spec_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/ncu_trace_spec.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:'gather' -s 2 -c 2 -f -o $O/gather_full \
    python tools/bench_photon.py > $O/ncu_gather.log 2>&1
ls -la $O
