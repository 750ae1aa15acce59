# ncu --set full on the first N k_isect3 launches of a frame (serial streams); plain run first.  usage: gpu_ncu_isect.sh [count] [name]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-14}; NAME=${2:-prof_isect}
B200RT_NO_CONCURRENCY=1 python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/plain.log 2>&1 &&
B200RT_NO_CONCURRENCY=1 ncu --set full --clock-control none --import-source on -k regex:"k_isect3" -c $N -o gpurun_out/$NAME python bench.py --steps 1 --warmup 1 --no-cpu > gpurun_out/ncu_isect.log 2>&1
tail -2 gpurun_out/ncu_isect.log | cut -c1-200
