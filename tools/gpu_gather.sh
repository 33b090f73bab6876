#!/bin/bash
# gather / photon variants: parity tests, stage timings, ncu of the gather + photon kernels
TAG=${1:-g}
O=gpurun_out/$TAG
mkdir -p $O
B2RT_GATHER_VARIANT=1 B2RT_PHOTON_VARIANT=1 timeout 900 python -m pytest tests -m gpu -x -q -k "photon or render or gather or dropin" > $O/pytest_v1.log 2>&1; echo "rc=$?" >> $O/pytest_v1.log
B2RT_GATHER_VARIANT=0 B2RT_PHOTON_VARIANT=0 timeout 300 python tools/bench_photon.py > $O/photon_g0p0.log 2>&1
B2RT_GATHER_VARIANT=1 B2RT_PHOTON_VARIANT=1 timeout 300 python tools/bench_photon.py > $O/photon_g1p1.log 2>&1
NCU="ncu --clock-control none"
B2RT_GATHER_VARIANT=1 B2RT_PHOTON_VARIANT=1 timeout 900 $NCU --set full --import-source on -k regex:'gather|photon_trace_regen' -s 20 -c 6 -f -o $O/gather_full \
    python tools/bench_photon.py > $O/ncu_gather.log 2>&1
ls -la $O
