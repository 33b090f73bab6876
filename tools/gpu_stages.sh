#!/bin/bash
# full GPU suite + roofline fractions of the HBM-bound stages + ncu of the sort and gather kernels
TAG=${1:-s}
O=gpurun_out/$TAG
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
timeout 900 python tools/bench_stages.py > $O/stages.json 2> $O/stages.err; echo "rc=$?" >> $O/stages.err
NCU="ncu --clock-control none"
timeout 900 $NCU --set full --import-source on -k regex:'radix_scatter|radix_hist' -s 4 -c 4 -f -o $O/sort_full \
    python tools/bench_stages.py 24 > $O/ncu_sort.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:'gather_tiled' -c 2 -f -o $O/gather_full \
    python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_gather.log 2>&1
timeout 300 python tools/bench_photon.py > $O/photon.log 2>&1
ls -la $O
