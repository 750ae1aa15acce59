# sweep an env var over values on one workload: bash tools/gpu_sweep.sh VAR "v1 v2 ..." [workload] [steps] [emulate_world]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
VAR=$1; VALS=$2; WL=${3:-c4}; ST=${4:-5}; EMU=${5:-0}
for v in $VALS; do
  env $VAR=$v B200RT_EMULATE_WORLD=$EMU timeout 300 python bench.py --workload $WL --steps $ST --warmup 3 --no-cpu > gpurun_out/sw.json 2> gpurun_out/sw.err
  python - "$VAR=$v" <<'PY'
import json,sys
try:
    d=json.loads(open('gpurun_out/sw.json').read().strip().splitlines()[-1]); r=d['roofline']
    print(sys.argv[1], 'ms', round(d['ms_per_step'],3), 'closest', round(r['kernel_ms_closest'],3), 'shadow', round(r['kernel_ms_shadow'],3), 'other', round(r['other_ms'],3), 'e2e', round(d['e2e']['ms_per_step'],3))
except Exception as e: print(sys.argv[1], 'ERR', e, open('gpurun_out/sw.err').read()[-600:])
PY
done
