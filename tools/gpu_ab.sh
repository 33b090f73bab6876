#!/bin/bash
# A/B pass: kernel variants (same results, different schedules), parity tests under each, ncu of the gather kernels.
TAG=${1:-ab}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_default.log 2>&1; echo "rc=$?" >> $O/pytest_default.log
B2RT_TRACE_VARIANT=1 B2RT_TRACE_REFILL=16 B2RT_GATHER_VARIANT=0 timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_alt.log 2>&1; echo "rc=$?" >> $O/pytest_alt.log
{
python tools/ab_trace.py
for rf in 0 8 16 24 32; do B2RT_TRACE_VARIANT=1 B2RT_TRACE_REFILL=$rf python tools/ab_trace.py; done
} > $O/ab_trace.log 2>&1
B2RT_GATHER_VARIANT=0 timeout 300 python tools/bench_photon.py > $O/photon_g0.log 2>&1
B2RT_GATHER_VARIANT=1 timeout 300 python tools/bench_photon.py > $O/photon_g1.log 2>&1
timeout 600 python bench.py --steps 50 > $O/bench.json 2> $O/bench.err
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:'gather' -c 4 -f -o $O/gather_full \
    python tools/bench_photon.py > $O/ncu_gather.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:'render_li_kernel|grid_reorder_kernel|grid_keys_kernel' -c 6 -f -o $O/render_full \
    python tools/bench_photon.py > $O/ncu_render.log 2>&1
ls -la $O
