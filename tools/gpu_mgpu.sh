#!/bin/bash
# N-GPU pass: parity tests on GPU 0, multi-GPU identity check, bench at N ranks (torchrun, NCCL)
TAG=${1:-m}; N=${2:-2}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?" >> $O/pytest_gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 tools/mgpu_check.py 24000000 1920 1080 > $O/mgpu_check.log 2>&1; echo "rc=$?" >> $O/mgpu_check.log
timeout 900 $TR --master-port 29512 bench.py --gpus $N --steps 100 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$?" >> $O/bench_n$N.err
timeout 600 $TR --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/bench_ref_n$N.json 2> $O/bench_ref_n$N.err
timeout 600 python bench.py --steps 100 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err
ls -la $O
