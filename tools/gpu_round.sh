#!/bin/bash
# One GPU-box pass: parity tests, bench (both arms), photon-frame stage timings, ncu launch lists and full captures.
# usage: tools/gpu_round.sh <tag>      (outputs under gpurun_out/<tag>/)
TAG=${1:-r01}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi > $O/smi.txt 2>&1
nproc > $O/nproc.txt
if [ -z "$SKIP_TESTS" ]; then
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
fi
timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err
timeout 600 python tools/bench_photon.py > $O/photon.log 2>&1
timeout 900 python tools/bench_cfg4.py > $O/cfg4.json 2> $O/cfg4.err
if [ -z "$SKIP_NCU" ]; then
NCU="ncu --clock-control none"
timeout 900 $NCU --metrics gpu__time_duration.sum -c 400 --csv --log-file $O/launches_bench.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/ncu_bench.log 2>&1
timeout 900 $NCU --metrics gpu__time_duration.sum -c 2000 --csv --log-file $O/launches_photon.csv \
    python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_photon.log 2>&1
timeout 900 $NCU --set full --import-source on -k regex:trace_ -s 16 -c 2 -f -o $O/trace_full \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/ncu_trace_full.log 2>&1
timeout 1200 $NCU --set full --import-source on -k regex:'gather_block|photon_trace_regen|render_li_kernel|grid_reorder_kernel|grid_keys_kernel' -s 4 \
    -c 14 -f -o $O/photon_full python tools/bench_photon.py 8000000 2000000 1920 1080 64 1 > $O/ncu_photon_full.log 2>&1
fi
ls -la $O
