#!/bin/bash
# launch list of one C1 / C2 frame (small, launch-bound configs)
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
for wl in c1 c2; do
python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu > gpurun_out/small_plain.log 2>&1 &&
B200RT_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_${wl}_launches.csv python bench.py --workload $wl --steps 3 --warmup 2 --no-cpu > gpurun_out/small_ncu.log 2>&1
python - $wl <<'PY'
import csv,sys,re,collections
wl=sys.argv[1]
rows=[r for r in csv.DictReader(l for l in open(f'gpurun_out/r02_{wl}_launches.csv') if not l.startswith('=='))]
# last full frame: from the last k_set_cam backwards
names=[(re.search(r'(k_\w+)',r['Kernel Name']).group(1) if re.search(r'(k_\w+)',r['Kernel Name']) else r['Kernel Name'][:20], float(r['Metric Value'].replace(',',''))/ (1000.0 if r['Metric Unit'] in ('ns','nsecond') else 1.0)) for r in rows]
idx=[i for i,(n,t) in enumerate(names) if n=='k_set_cam']
fr=names[idx[-3]:idx[-2]]
print(wl, 'kernels in frame', len(fr), 'sum us', round(sum(t for n,t in fr),1))
agg=collections.OrderedDict()
for n,t in fr: agg.setdefault(n,[0,0.0]); agg[n][0]+=1; agg[n][1]+=t
print({k:(v[0],round(v[1],1)) for k,v in agg.items()})
PY
done
