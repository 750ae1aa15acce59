# C4 on 8 GPUs (torchrun) under scheduling variants.  usage: bash tools/gpu_n8_variants.sh
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
i=0
for k in "X=1" "B200RT_GRAPH_ALL=1" "B200RT_NO_CONCURRENCY=1" "X=2"; do
  i=$((i+1))
  env $k timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2960$i bench.py --gpus 8 --steps 20 --warmup 3 --no-cpu --no-dense > gpurun_out/n8_var$i.json 2> gpurun_out/n8_var$i.err
  python -c "
import json
j=json.loads(open('gpurun_out/n8_var$i.json').read().strip().splitlines()[-1])
print('$k', 'ms', round(j['ms_per_step'],4), 'e2e', round(j['e2e']['ms_per_step'],4), 'enq', round(j['host_enqueue_ms_per_step'],3), j['e2e'].get('checksum_matches_one_gpu_frame'))"
done
