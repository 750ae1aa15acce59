# final numbers of a build: gpu tests, smoke, benches of all configs + reference arm, ncu launch list and per-launch metrics.  usage: gpu_final.sh TAG
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out; T=${1:-v11}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/${T}_pytest_gpu.log; tail -1 gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
for wl in c1 c2 c3 c5; do timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu > gpurun_out/${T}_bench_$wl.json 2> gpurun_out/${T}_bench_$wl.err; done
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
python bench.py --steps 2 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_ncu1.log 2>&1
python bench.py --steps 1 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_plain3.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:"k_isect3|k_setup|k_shade|k_shadow_shade|k_key|k_perm|k_raygen|k_resolve|k_sgrid|k_mrank|k_clus" -c 84 --csv --log-file gpurun_out/${T}_frame.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_ncu3.log 2>&1
tail -c 300 gpurun_out/${T}_bench_c4.json
