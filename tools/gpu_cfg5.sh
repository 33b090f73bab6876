#!/bin/bash
# Cfg5 photon sweep at N ranks (N=1: plain python); each point in its own process and timeout
TAG=${1:-c5}; N=${2:-1}
O=gpurun_out/$TAG
mkdir -p $O
if [ "$N" = "1" ]; then RUN="python"; else RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521"; fi
for P in 2e6 2e7 2e8; do
  timeout 600 $RUN tools/bench_cfg5.py $P >> $O/cfg5_n$N.jsonl 2>> $O/cfg5_n$N.err; echo "paths=$P rc=$?" >> $O/cfg5_n$N.err
done
if [ -n "$BIG" ]; then
  timeout 900 $RUN tools/bench_cfg5.py $BIG >> $O/cfg5_n$N.jsonl 2>> $O/cfg5_n$N.err; echo "paths=$BIG rc=$?" >> $O/cfg5_n$N.err
fi
if [ "$N" != "1" ]; then
  timeout 900 $RUN tools/bench_cfg4.py > $O/cfg4_n$N.json 2> $O/cfg4_n$N.err; echo "rc=$?" >> $O/cfg4_n$N.err
fi
nvidia-smi --query-gpu=memory.used,memory.total --format=csv > $O/mem_after.txt
ls -la $O
