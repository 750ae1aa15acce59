# ncu --set full on the per-hit kernels (k_setup, k_shade, k_raygen) of the first frame; plain run first
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
B200RT_NO_CONCURRENCY=1 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
B200RT_NO_CONCURRENCY=1 ncu --set full --clock-control none --import-source on -k regex:"k_setup" -c 2 -o gpurun_out/prof_setup python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_small.log 2>&1
tail -2 gpurun_out/ncu_small.log | cut -c1-200
