#!/bin/bash
# interleaved image stripes at N ranks: identity check + Cfg4 frame with 1 / 2 / 4 / 8 stripes per rank
TAG=${1:-stripes}; N=${2:-2}
O=gpurun_out/$TAG
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 tools/mgpu_check.py 24000000 1920 1080 > $O/mgpu_check.log 2>&1; echo "rc=$?" >> $O/mgpu_check.log
grep -h "identical\|rep 2" $O/mgpu_check.log
for s in 1 2 4 8; do
  STRIPES=$s timeout 400 $TR --master-port 2952$s tools/bench_cfg4.py 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('stripes $s: cfg4 ms_per_frame %.2f all %s' % (d['ms_per_frame'], [round(v,1) for v in d['all_ms']]))" | tee -a $O/cfg4_stripes.txt
done
