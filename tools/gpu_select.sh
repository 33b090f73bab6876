#!/bin/bash
# threshold-selection gather: parity tests (all gather schedules), stage timings, ncu of the block gather kernel
TAG=${1:-sel}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q -k "photon or render or gather or dropin or variants" > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
timeout 300 python tools/bench_photon.py > $O/photon.log 2>&1
B2RT_GATHER_VARIANT=1 timeout 300 python tools/bench_photon.py > $O/photon_g1.log 2>&1
if [ -z "$SKIP_NCU" ]; then
timeout 900 ncu --clock-control none --set full --import-source on -k regex:'gather_block' -c 2 -f -o $O/gather_full \
    python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_gather.log 2>&1
fi
ls -la $O
