# r01 v7: gpu tests, smoke, benches of all configs, then the ncu passes (each after a plain run of the same command exited 0)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
bash tools/gpu_try.sh
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/v7_bench_full.json 2> gpurun_out/v7_bench_full.err
python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/v7_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/v7_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/v7_ncu1.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v7_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:"k_isect3|k_setup|k_shade|k_regroup|k_raygen" -c 70 --csv --log-file gpurun_out/v7_frame.csv python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v7_ncu3.log 2>&1
B200RT_NO_CONCURRENCY=1 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v7_plain2.log 2>&1 &&
B200RT_NO_CONCURRENCY=1 ncu --set full --clock-control none --import-source on -k regex:"k_isect3|k_setup" -c 18 -o gpurun_out/v7_prof python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/v7_ncu2.log 2>&1
tail -2 gpurun_out/v7_ncu2.log | cut -c1-200; ls -la gpurun_out | tail -20
