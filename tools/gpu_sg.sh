cd $GRAFT_REPO_ROOT
python bench.py --steps 2 --warmup 1 --no-cpu 2>/dev/null | tail -1 > gpurun_out/sg_bench.json
python -c "
import json; d=json.load(open('gpurun_out/sg_bench.json')); r=d['roofline']
print('ms', d['ms_per_step'], 'pivot tests', r['executed_tests_pivot_form'], 'general', r['executed_tests_general_form'], 'exact/ray', r['exact_tests_per_ray'], 'traced frac', r['shadow_queries_traced_fraction'])"
B200RT_NO_CONCURRENCY=1 ncu --set full --clock-control none --import-source on -k regex:"k_shadow_grid" --launch-skip 8 -c 4 -o gpurun_out/sg_prof python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/sg_ncu.log 2>&1
tail -1 gpurun_out/sg_ncu.log | cut -c1-100
