cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for np in 4; do
  B200RT_NP=$np timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/bench_np$np.json 2> gpurun_out/bench_np$np.err
done
tail -3 gpurun_out/pytest_gpu.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_np*.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        r=d['roofline']
        print(f, d['ms_per_step'], d['value'], r['frac'], r['kernel_ms_per_frame'], r.get('kernel_ms_closest'), r.get('kernel_ms_shadow'), r.get('other_ms'), r['slow_path_entries_per_ray'], d['e2e']['ms_per_step'])
    except Exception as e: print(f, 'ERR', e)
PY
