# bench one workload under several environment settings.  usage: gpu_env_sweep.sh workload steps "ENV1=a ENV2=b" "ENV3=c" ...
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
WL=$1; ST=$2; shift 2
for e in "" "$@"; do
  timeout 300 env $e python bench.py --workload $WL --steps $ST --warmup 3 --no-cpu > gpurun_out/sweep.json 2> gpurun_out/sweep.err
  python - "$e" <<'PY'
import json,sys
try:
    d=json.loads(open('gpurun_out/sweep.json').read().strip().splitlines()[-1]); r=d['roofline']
    print('[%s]'%sys.argv[1], 'ms', round(d['ms_per_step'],3), 'host', round(d.get('host_enqueue_ms_per_step',0),3), 'closest', round(r['kernel_ms_closest'],3), 'shadow', round(r['kernel_ms_shadow'],3), 'other', round(r['other_ms'],3), 'e2e', round(d['e2e']['ms_per_step'],3))
except Exception as ex: print(sys.argv[1], 'ERR', ex, open('gpurun_out/sweep.err').read()[-400:])
PY
done
