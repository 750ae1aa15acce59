# ncu launch list of the production bench command (no dense-mode frames): usage gpu_launchlist.sh TAG
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out; T=${1:-v9}
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 2 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/${T}_launches.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-dense > gpurun_out/${T}_ncu1.log 2>&1
tail -1 gpurun_out/${T}_plain.log | cut -c1-120
