#!/bin/bash
# launch list (device time of every kernel) of a short bench.py run on one GPU -- after the same command exited 0 without ncu
cd "$(dirname "${BASH_SOURCE[0]}")/.." || exit 1
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 6 --warmup 3 --no-verify --no-baseline"
$CMD > $OUT/ll_plain.json 2> $OUT/ll_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_bench.csv $CMD > $OUT/ll_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/launches_bench.csv")) if len(r) > 5]
hdr = rows[0]
k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
u = hdr.index("Metric Unit")
agg = collections.OrderedDict()
for r in rows[1:]:
    try:
        t = float(r[v].replace(",", ""))
    except ValueError:
        continue
    t = t / 1e3 if r[u] in ("ns", "nsecond") else t
    a = agg.setdefault(r[k][:90], [0, 0.0])
    a[0] += 1
    a[1] += t
print("launches  total_us  kernel")
for name, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
    print(f"{n:8d} {t:9.1f}  {name}")
PY
exit 0
