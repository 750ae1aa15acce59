# Scaling of one C4 frame (column strips) and of the C5 camera-pose batch (frames) over the GPUs of one box.  usage: bash tools/gpu_scale.sh [NG] [TAG]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NG=${1:-8}; T=${2:-r02}
run() { # workload n steps
  if [ $2 -eq 1 ]; then timeout 400 python bench.py --workload $1 --steps $3 --warmup 3 --no-cpu --no-dense > gpurun_out/${T}_scale_$1_n$2.json 2> gpurun_out/${T}_scale_$1_n$2.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $2 --master-addr 127.0.0.1 --master-port 2950$2 bench.py --workload $1 --gpus $2 --steps $3 --warmup 3 --no-cpu --no-dense > gpurun_out/${T}_scale_$1_n$2.json 2> gpurun_out/${T}_scale_$1_n$2.err; fi
}
for n in 1 2 4 8; do if [ $n -le $NG ]; then run c4 $n 20; fi; done
for n in 1 2 4 8; do if [ $n -le $NG ]; then run c5 $n 64; fi; done
python - $T <<'PY'
import json,glob,sys
T=sys.argv[1]
for wl in ("c4","c5"):
    base=None
    for n in (1,2,4,8):
        f='gpurun_out/%s_scale_%s_n%d.json'%(T,wl,n)
        try:
            d=json.loads(open(f).read().strip().splitlines()[-1])
            if n==1: base=d['ms_per_step']; be=d['e2e']['ms_per_step']
            print(wl, n, "ms", round(d["ms_per_step"],3), "x", round(base/d["ms_per_step"],2), "e2e ms", round(d["e2e"]["ms_per_step"],3), "x", round(be/d["e2e"]["ms_per_step"],2), "assembled", d["e2e"].get("assembled"), "checksum ok", d["e2e"].get("checksum_matches_one_gpu_frame"), d["clocks"]["sm_mhz"], d["clocks"]["reasons"])
        except Exception as e: print(f,'ERR',e, open(f.replace('.json','.err')).read()[-800:] if glob.glob(f.replace('.json','.err')) else '')
PY
