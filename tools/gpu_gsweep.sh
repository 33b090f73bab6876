#!/bin/bash
# block gather: first-radius target x shared-tile capacity sweep (camera pass time of tools/bench_photon.py, last rep)
TAG=${1:-gsweep}
O=gpurun_out/$TAG
mkdir -p $O
for tcap in ${TCAPS:-2048 1536 1024}; do for tg in ${TARGETS:-1.3 1.6 2.0 2.5}; do
  echo -n "tcap $tcap target $tg: " >> $O/sweep.txt
  B2RT_GATHER_TCAP=$tcap B2RT_GATHER_TARGET=$tg timeout 300 python tools/bench_photon.py 2>&1 | tail -1 | grep -o "frame [0-9.]* ms | direct" >> $O/sweep.txt || echo fail >> $O/sweep.txt
done; done
cat $O/sweep.txt
