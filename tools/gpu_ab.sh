# A/B of prebuilt libraries under tools/ab/*.so on one workload: bash tools/gpu_ab.sh [workload] [steps] [emulate_world]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
WL=${1:-c4}; ST=${2:-10}
cp python-ray-tracer_b200/b200rt/libb200rt.so /tmp/orig.so
for f in tools/ab/*.so; do
  n=$(basename $f .so)
  cp $f python-ray-tracer_b200/b200rt/libb200rt.so; touch python-ray-tracer_b200/b200rt/libb200rt.so
  timeout 300 env B200RT_EMULATE_WORLD=${3:-0} python bench.py --workload $WL --steps $ST --warmup 3 --no-cpu --no-dense > gpurun_out/ab_$n.json 2> gpurun_out/ab_$n.err
  python - $n <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open('gpurun_out/ab_%s.json'%n).read().strip().splitlines()[-1]); r=d['roofline']['exact_culled']['kernel_ms']
    print(n, "ms", round(d["ms_per_step"],3), "host", round(d.get("host_enqueue_ms_per_step",0),3), 'closest', round(r['closest_hit'],3), 'shadow', round(r['shadow_shade'],3), 'other', round(r['other'],3), 'e2e', round(d['e2e']['ms_per_step'],3), flush=True)
except Exception as e: print(n, 'ERR', e, open('gpurun_out/ab_%s.err'%n).read()[-600:])
PY
done
cp /tmp/orig.so python-ray-tracer_b200/b200rt/libb200rt.so
