cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NG=${1:-8}
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --no-dense > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
for n in 2 4 8; do
  if [ $n -le $NG ]; then
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 10 --warmup 3 --no-cpu --no-dense > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
  fi
done
python - <<'PY'
import json,glob
base=None
for n in (1,2,4,8):
    f='gpurun_out/scale_n%d.json'%n
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1])
        if n==1: base=d['ms_per_step']; be=d['e2e']['ms_per_step']
        print(n, "ms", round(d["ms_per_step"],3), "x", round(base/d["ms_per_step"],2), "e2e ms", round(d["e2e"]["ms_per_step"],3), "x", round(be/d["e2e"]["ms_per_step"],2), "assembled", d["e2e"].get("assembled"), "checksum ok", d["e2e"].get("checksum_matches_one_gpu_frame"), d["clocks"])
    except Exception as e: print(f,'ERR',e, open(f.replace('.json','.err')).read()[-1500:] if glob.glob(f.replace('.json','.err')) else '')
PY
