"""Per-source-line instruction / stall-sample shares from `ncu -i rep --page source --csv --print-source cuda,sass`.
python tools/ncu_lines.py src.csv [top]"""
import collections
import csv
import sys


def main(path, top=40):
    rows = list(csv.reader(open(path)))
    cur, hdr, ix = None, None, {}
    agg, smp, thr, src = collections.Counter(), collections.Counter(), collections.Counter(), {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur, hdr = r[1], None
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            ix = {k: i for i, k in enumerate(hdr)}
            continue
        if hdr is None:
            continue
        try:
            ln = int(r[0])
        except ValueError:
            continue
        key = (cur.split("/")[-1], ln)

        def num(col):
            try:
                return int(r[ix[col]])
            except (ValueError, KeyError, IndexError):
                return 0
        agg[key] += num("Instructions Executed")
        thr[key] += num("Thread Instructions Executed")
        smp[key] += num("# Samples")
        src[key] = r[1].strip()[:100]
    tot, ts = sum(agg.values()), max(1, sum(smp.values()))
    print("total warp instructions %d, thread instructions %d (%.1f threads/inst), %d stall samples" %
          (tot, sum(thr.values()), sum(thr.values()) / max(1, tot), ts))
    print("inst%  smp%   thr/inst  line")
    for k, v in agg.most_common(top):
        print("%5.1f %5.1f  %5.1f  %s:%d  %s" % (100 * v / tot, 100 * smp[k] / ts, thr[k] / max(1, v), k[0], k[1], src[k]))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
