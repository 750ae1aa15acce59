cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_setup|k_shade|k_raygen|k_resolve" -c 60 --csv --log-file gpurun_out/small_launches.csv python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_list.log 2>&1
python - <<'PY'
import csv,collections,re
rows=[r for r in csv.reader(open('gpurun_out/small_launches.csv')) if len(r)>5]
hdr=[i for i,r in enumerate(rows) if r[0]=='ID'][0]
h=rows[hdr]; data=rows[hdr+1:]
iK=h.index('Kernel Name'); iV=h.index('Metric Value')
for r in data[:22]: print(r[iK][:20], float(r[iV].replace(',',''))/1e6)
PY
