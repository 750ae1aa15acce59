cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
N=$1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2970$N bench.py --gpus $N --steps 20 --warmup 3 --no-cpu --no-dense > gpurun_out/r02f_scale_c4_n${N}_final.json 2> gpurun_out/n${N}.err
python -c "
import json
j=json.loads(open('gpurun_out/r02f_scale_c4_n${N}_final.json').read().strip().splitlines()[-1])
print($N, 'ms', round(j['ms_per_step'],4), 'e2e', round(j['e2e']['ms_per_step'],4), j['e2e'].get('checksum_matches_one_gpu_frame'))"
