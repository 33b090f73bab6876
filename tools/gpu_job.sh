#!/bin/bash
# One parametrised GPU-box job list (replaces the per-experiment scripts of round 1).
# usage: tools/gpu_job.sh <tag> <job> [<job> ...]       outputs under gpurun_out/<tag>/
# jobs: tests | tests:<pytest -k expr> | fgtests | bench | benchref | fg540 | fg1080 | photon | stages | cfg4 | smoke
TAG=$1; shift
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/smi.txt 2>&1
nproc > $O/nproc.txt
for job in "$@"; do
  case $job in
    tests)    timeout 1700 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log ;;
    tests:*)  timeout 900 python -m pytest tests -m gpu -x -q -k "${job#tests:}" > $O/pytest_k.log 2>&1; echo "pytest rc=$?" >> $O/pytest_k.log ;;
    multitests) timeout 900 python -m pytest tests/test_gpu_multi_capi.py -m gpu -x -q > $O/pytest_multi.log 2>&1; echo "pytest rc=$?" >> $O/pytest_multi.log ;;
    fgtests)  timeout 600 python -m pytest tests/test_gpu_zfinalgather.py tests/test_gpu_photon.py -m gpu -x -q > $O/pytest_fg.log 2>&1; echo "pytest rc=$?" >> $O/pytest_fg.log ;;
    smoke)    timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log ;;
    bench)    timeout 900 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err ;;
    benchq)   timeout 600 python bench.py --steps 20 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err ;;
    bench2)   timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 50 > $O/bench_n2.json 2> $O/bench_n2.err; echo "rc=$?" >> $O/bench_n2.err ;;
    bench4)   timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 4 --steps 50 > $O/bench_n4.json 2> $O/bench_n4.err; echo "rc=$?" >> $O/bench_n4.err ;;
    bench8)   timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus 8 --steps 50 > $O/bench_n8.json 2> $O/bench_n8.err; echo "rc=$?" >> $O/bench_n8.err ;;
    ncufg)    python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/plain_fg.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'fg_lookup_kernel|fg_trace_kernel|fg_combine_kernel' -c 3 -f -o $O/fg_split python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/ncu_fg.log 2>&1 ;;
    ncutrace) python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/plain_bench.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:trace_spec -s 8 -c 2 -f -o $O/trace_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-photon-extra > $O/ncu_trace.log 2>&1 ;;
    launchbench) python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/plain_bench_short.json 2> $O/plain_bench_short.err && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_bench.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/ncu_launches_bench.log 2>&1 ;;
    instances) timeout 500 python tools/bench_instances.py 1000 20000 > $O/bench_instances_1000x20k.jsonl 2> $O/bench_instances.err ;;
    tiles:*)  timeout 600 python tools/bench_tiles.py ${job#tiles:} > $O/tiles_n${job#tiles:}.jsonl 2> $O/tiles.err ;;
    sanmem)   python -c "import __graft_entry__ as g; g.smoke()" > $O/plain_smoke.log 2>&1 && timeout 1500 compute-sanitizer --tool memcheck --log-file $O/sanitizer_memcheck.txt python -c "import __graft_entry__ as g; g.smoke()" > $O/san_memcheck_run.log 2>&1; echo "rc=$?" >> $O/san_memcheck_run.log ;;
    sanrace)  python -c "import __graft_entry__ as g; g.smoke()" > $O/plain_smoke.log 2>&1 && timeout 1500 compute-sanitizer --tool racecheck --log-file $O/sanitizer_racecheck.txt python -c "import __graft_entry__ as g; g.smoke()" > $O/san_racecheck_run.log 2>&1; echo "rc=$?" >> $O/san_racecheck_run.log ;;
    saninit)  python -c "import __graft_entry__ as g; g.smoke()" > $O/plain_smoke.log 2>&1 && timeout 1500 compute-sanitizer --tool initcheck --log-file $O/sanitizer_initcheck.txt python -c "import __graft_entry__ as g; g.smoke()" > $O/san_initcheck_run.log 2>&1; echo "rc=$?" >> $O/san_initcheck_run.log ;;
    exch:*)   timeout 600 python tools/bench_exchange.py ${job#exch:} > $O/exchange_n${job#exch:}.jsonl 2> $O/exchange.err ;;
    topo)     nvidia-smi topo -m > $O/topo.txt 2>&1 ;;
    abplanes) for v in 0 1 0 1; do B2RT_TRACE_PLANES=$v timeout 300 python bench.py --steps 100 --no-cpu-baseline --no-photon-extra > $O/bench_planes$v.json 2>> $O/abplanes.err; python -c "import json;d=json.load(open('$O/bench_planes$v.json'));print('planes',$v,d['value'],d['closest_hit_mrays'],d['any_hit_mrays'],d['incoherent_closest_hit_mrays'])" >> $O/abplanes.txt; done ;;
    tracestats) timeout 300 python tools/trace_stats.py > $O/trace_stats.json 2> $O/trace_stats.err ;;
    benchref) timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err ;;
    fg540)    timeout 300 python tools/bench_fg.py 1000000 100000 960 540 32 2 > $O/bench_fg_540.jsonl 2> $O/bench_fg_540.err; echo "rc=$?" >> $O/bench_fg_540.err ;;
    fg540v0)  B2RT_FG_VARIANT=0 timeout 300 python tools/bench_fg.py 1000000 100000 960 540 32 1 > $O/bench_fg_540_v0.jsonl 2> $O/bench_fg_540_v0.err ;;
    fg1080)   timeout 600 python tools/bench_fg.py 8000000 2000000 1920 1080 32 2 > $O/bench_fg_1080.jsonl 2> $O/bench_fg_1080.err; echo "rc=$?" >> $O/bench_fg_1080.err ;;
    photon)   timeout 600 python tools/bench_photon.py > $O/photon.log 2>&1 ;;
    converged) timeout 900 python tools/converged_parity.py --gpu > $O/converged.jsonl 2> $O/converged.err ;;
    stages)   timeout 600 python tools/bench_stages.py > $O/stages.json 2> $O/stages.err ;;
    buildlist) python tools/bench_stages.py 20 > $O/plain_stages.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_stages.csv python tools/bench_stages.py 20 > $O/ncu_stages.log 2>&1 ;;
    stages0)  B2RT_SORT_VARIANT=0 timeout 600 python tools/bench_stages.py > $O/stages_sort_variant0.json 2> $O/stages0.err ;;
    cfg4)     timeout 900 python tools/bench_cfg4.py > $O/cfg4.json 2> $O/cfg4.err ;;
    *)        echo "unknown job $job" >> $O/jobs.err ;;
  esac
done
ls -la $O
