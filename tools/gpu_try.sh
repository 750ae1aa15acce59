cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 > gpurun_out/pytest_gpu.log; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_nr8.json 2> gpurun_out/bench_nr8.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_nr8.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f, d['ms_per_step'], d['value'], r['frac'], r['kernel_ms_per_frame'], r.get('kernel_ms_closest'), r.get('kernel_ms_shadow'), r.get('other_ms'), r['slow_path_entries_per_ray'], r.get('exact_tests_per_ray'), r.get('shadow_queries_traced_fraction'), d['e2e']['ms_per_step'])
    except Exception as e: print(f, 'ERR', e, open(f.replace('.json','.err')).read()[-800:])
PY
