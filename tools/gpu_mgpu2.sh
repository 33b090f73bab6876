#!/bin/bash
TAG=${1:-m2}; N=${2:-2}
O=gpurun_out/$TAG
mkdir -p $O
bash tools/gpu_mgpu.sh $TAG $N > $O/mgpu_sh.log 2>&1
bash tools/gpu_cfg5.sh $TAG $N > $O/cfg5_sh.log 2>&1
{ for ch in 65536 131072 262144 524288; do B2RT_HOST_CHUNK=$ch timeout 300 python tools/ab_e2e.py; done; } > $O/ab_e2e.log 2>&1
ls -la $O
