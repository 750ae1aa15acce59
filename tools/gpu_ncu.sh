# r01 v5 profiles: launch list of one bench command, full capture of k_isect3 launches of the first frame, dram traffic per launch
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/v6_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/v6_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/v6_ncu1.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v6_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_isect3 -c 14 -o gpurun_out/v6_prof_isect python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v6_ncu2.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v6_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:k_isect3 -c 40 --csv --log-file gpurun_out/v6_isect_frame.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v6_ncu3.log 2>&1
tail -2 gpurun_out/v6_ncu2.log; ls -la gpurun_out | grep v5
