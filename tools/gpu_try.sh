cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/pytest_gpu.log; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for wl in c1 c2 c3 c5 c4; do
timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_c[1-5].json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f, d['config']['workload'], 'ms', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],1), 'frac', round(r['frac'],3), 'isect', round(r['kernel_ms_per_frame'],3), 'other', round(r['other_ms'],3), 'e2e ms', round(d['e2e']['ms_per_step'],3), 'rays/px', round(d['config']['rays_per_pixel'],2))
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-800:])
PY
