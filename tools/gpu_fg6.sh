#!/bin/bash
# fully warp-cooperative lookups (B2RT_FG_VARIANT=2): parity tests + fg frame timing
O=gpurun_out/${1:-r03h}
mkdir -p $O
B2RT_TEST_FG_VARIANTS=0,1,2 timeout 25 python -m pytest tests/test_gpu_zfinalgather.py -m gpu -q -x > $O/pytest_fg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_fg.log
B2RT_FG_VARIANT=2 timeout 12 python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/bench_fg_v2.jsonl 2> $O/bench_fg_v2.err; echo "rc=$?" >> $O/bench_fg_v2.err
