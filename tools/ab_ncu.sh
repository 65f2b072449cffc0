#!/bin/bash
# tools/ab_ncu.sh <kernel regex> <profile_once workload> — duration and DRAM bytes of the selected launches for the in-tree library
# and every alt_builds/*/libgoofit_b200.so, one ncu pass each (three metrics, so replays are cheap). Not a benchmark.
rx=$1; kind=$2
cp goofit_b200/lib/libgoofit_b200.so /tmp/new.so
for v in /tmp/new.so alt_builds/*/libgoofit_b200.so; do
  cp $v goofit_b200/lib/libgoofit_b200.so
  timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:$rx --csv \
      --log-file /tmp/ab_ncu.csv python tools/profile_once.py 10000000 $kind > /tmp/ab_ncu.log 2>&1 || tail -3 /tmp/ab_ncu.log
  python - "$v" <<'P'
import csv, sys
rows = list(csv.reader(open('/tmp/ab_ncu.csv')))
hi = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
h = rows[hi]
k, m, val, unit = h.index('Kernel Name'), h.index('Metric Name'), h.index('Metric Value'), h.index('Metric Unit')
per = {}
for r in rows[hi + 1:]:
    per.setdefault((int(r[0]), r[k].split('(')[0].replace('void ', '')), {})[r[m]] = (float(r[val].replace(',', '')), r[unit])
out = []
for (i, name), d in sorted(per.items()):
    t, tu = d['gpu__time_duration.sum']
    t = t / 1e3 if tu in ('ns', 'nsecond') else (t if tu in ('us', 'usecond') else t * 1e3)
    f = lambda x: x[0] * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[x[1]] / 1e9
    out.append('%s %.0fus r%.2f w%.2f' % (name[-3:], t, f(d['dram__bytes_read.sum']), f(d['dram__bytes_write.sum'])))
print(sys.argv[1].split('/')[-2], ' | '.join(out))
P
done
cp /tmp/new.so goofit_b200/lib/libgoofit_b200.so
