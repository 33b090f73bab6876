#!/bin/bash
# traversal variants A/B + gather re-check
TAG=${1:-t}
O=gpurun_out/$TAG
mkdir -p $O
{
python tools/ab_trace.py
for ch in 32 256 1024; do for rf in 0 16; do B2RT_TRACE_VARIANT=1 B2RT_TRACE_REFILL=$rf B2RT_TRACE_CHUNK=$ch python tools/ab_trace.py; echo "  ^ chunk=$ch"; done; done
} > $O/ab_trace.log 2>&1
timeout 300 python tools/bench_photon.py > $O/photon_default.log 2>&1
B2RT_TRACE_VARIANT=1 B2RT_TRACE_REFILL=16 timeout 600 python -m pytest tests -m gpu -x -q -k "parity" > $O/pytest_pt.log 2>&1; echo "rc=$?" >> $O/pytest_pt.log
NCU="ncu --clock-control none"
B2RT_TRACE_VARIANT=1 B2RT_TRACE_REFILL=16 timeout 900 $NCU --set full --import-source on -k regex:trace_persistent -s 8 -c 2 -f -o $O/trace_pt_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/ncu_trace_pt.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:'gather' -s 2 -c 2 -f -o $O/gather_full \
    python tools/bench_photon.py > $O/ncu_gather.log 2>&1
ls -la $O
