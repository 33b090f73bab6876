#!/bin/bash
TAG=${1:-g2}
O=gpurun_out/$TAG
mkdir -p $O
for qt in 8 16 32; do B2RT_GATHER_QT=$qt timeout 300 python tools/bench_photon.py > $O/photon_qt$qt.log 2>&1; echo "rc=$?" >> $O/photon_qt$qt.log; done
B2RT_GATHER_VARIANT=1 timeout 300 python tools/bench_photon.py > $O/photon_g1.log 2>&1
for qt in 8 16; do B2RT_GATHER_QT=$qt timeout 600 python -m pytest tests -m gpu -q -k "photon or gather or render or variants or dropin" > $O/pytest_qt$qt.log 2>&1; echo "rc=$?" >> $O/pytest_qt$qt.log; done
B2RT_GATHER_QT=8 timeout 600 ncu --clock-control none --set full --import-source on -k regex:'gather_block' -c 2 -f -o $O/gather_full \
    python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_gather.log 2>&1
ls -la $O
