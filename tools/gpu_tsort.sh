#!/bin/bash
# child ordering in trace_spec_kernel: full sort / nearest first / slot order; hashes must not change
TAG=${1:-tsort}
O=gpurun_out/$TAG
mkdir -p $O
for tris in 1000000 8000000; do for so in 2 1 0; do
  B2RT_TRACE_SORT=$so timeout 300 python tools/ab_trace.py $tris 2>&1 | tail -1 | sed "s/^/sort $so: /" >> $O/ab.txt
done; done
cat $O/ab.txt
