#!/bin/bash
# photon-map assembly A/B at N ranks: one padded ncclAllGather (default) vs one broadcast per rank
TAG=${1:-asm}; N=${2:-2}
O=gpurun_out/$TAG
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 tools/mgpu_check.py 24000000 1920 1080 > $O/mgpu_check.log 2>&1; echo "rc=$?" >> $O/mgpu_check.log
timeout 400 $TR --master-port 29514 tools/bench_cfg5.py 2e7 2e8 > $O/cfg5_allgather.jsonl 2> $O/cfg5_allgather.err; echo "rc=$?" >> $O/cfg5_allgather.err
B2RT_ASSEMBLE=broadcast timeout 400 $TR --master-port 29515 tools/bench_cfg5.py 2e7 2e8 > $O/cfg5_broadcast.jsonl 2> $O/cfg5_broadcast.err; echo "rc=$?" >> $O/cfg5_broadcast.err
grep -h "identical\|rep 2" $O/mgpu_check.log
grep -h '^{' $O/cfg5_allgather.jsonl $O/cfg5_broadcast.jsonl | python -c "
import sys, json
for l in sys.stdin:
    j = json.loads(l); print(j['n_gpus'], j['paths'], 'frame %.2f shoot %.2f assemble %.2f grid %.2f camera %.2f bus %s' % (j['frame_ms'], j['shoot_ms'], j['allgather_ms'], j['grid_ms'], j['camera_gather_ms'], j['allgather_bus_GBs']))"
