#!/bin/bash
# tight N-rank pass (charged N x): identities, bench, Cfg4, one Cfg5 point
TAG=${1:-n8}; N=${2:-8}
O=gpurun_out/$TAG
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 tools/mgpu_check.py 24000000 1920 1080 > $O/mgpu_check.log 2>&1; echo "rc=$?" >> $O/mgpu_check.log
timeout 400 $TR --master-port 29512 bench.py --gpus $N --steps 100 --warmup 3 --no-cpu-baseline > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?" >> $O/bench_n$N.err
timeout 400 $TR --master-port 29513 tools/bench_cfg4.py > $O/cfg4_n$N.json 2> $O/cfg4_n$N.err; echo "rc=$?" >> $O/cfg4_n$N.err
timeout 400 $TR --master-port 29514 tools/bench_cfg5.py 2e7 2e8 > $O/cfg5_n$N.jsonl 2> $O/cfg5_n$N.err; echo "rc=$?" >> $O/cfg5_n$N.err
ls -la $O
