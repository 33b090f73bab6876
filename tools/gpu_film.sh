#!/bin/bash
# film stage: roofline numbers of film_filter_kernel, filtered Cfg3 / Cfg4 frames, ncu of the film kernel
TAG=${1:-film}
O=gpurun_out/$TAG
mkdir -p $O
timeout 900 python tools/bench_stages.py 24 > $O/stages.json 2> $O/stages.err; echo "rc=$?" >> $O/stages.err
for f in "" mitchell; do
  FILTER=$f timeout 300 python tools/bench_photon.py 2>&1 | tail -1 | sed "s/^/filter '$f': /" >> $O/frames.txt
  FILTER=$f timeout 600 python tools/bench_cfg4.py 2>/dev/null | grep '^{' | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('filter \'$f\': cfg4 ms_per_frame %.2f all %s' % (d['ms_per_frame'], [round(v,1) for v in d['all_ms']]))" >> $O/frames.txt
done
cat $O/frames.txt
if [ -z "$SKIP_NCU" ]; then
FILTER=mitchell timeout 900 ncu --clock-control none --set full --import-source on -k regex:'film_' -c 2 -f -o $O/film_full \
    python tools/bench_cfg4.py > $O/ncu_film.log 2>&1
fi
ls -la $O
