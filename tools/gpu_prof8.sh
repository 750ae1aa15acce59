# r01 v8: ncu passes on the C4 bench (each after a plain run of the same command exited 0).  usage: gpu_prof8.sh [tag]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
T=${1:-v8}
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/${T}_ncu1.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/${T}_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:"k_isect3|k_setup|k_shade|k_key|k_perm|k_raygen|k_resolve" -c 100 --csv --log-file gpurun_out/${T}_frame.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/${T}_ncu3.log 2>&1
B200RT_NO_CONCURRENCY=1 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/${T}_plain2.log 2>&1 &&
B200RT_NO_CONCURRENCY=1 ncu --set full --clock-control none --import-source on -k regex:"k_isect3|k_setup" --launch-skip 36 -c 18 -o gpurun_out/${T}_prof python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/${T}_ncu2.log 2>&1
tail -2 gpurun_out/${T}_ncu2.log | cut -c1-200
