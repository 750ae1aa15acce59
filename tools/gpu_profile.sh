#!/bin/bash
# Profiling passes of a build (one ncu mode per gpurun call — /opt/skills/guides/B200_PROFILING.md).  usage: tools/gpu_profile.sh MODE [TAG]
#   launches   every launch of a short bench run with its device time                       -> gpurun_out/TAG_launches.csv
#   frame      per-launch metrics (time, DRAM bytes, FMA pipe, issue slots, instructions)   -> gpurun_out/TAG_frame.csv
#   dense      ncu --set full on the three k_isect3 launch types of the brute-force mode    -> gpurun_out/TAG_dense.ncu-rep
#   culled     ncu --set full on k_shadow_shade / k_closest_grid / k_setup / k_isect3<0>    -> gpurun_out/TAG_culled.ncu-rep
# The ncu command line is run plainly first (same arguments) and must exit 0.
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; MODE=${1:-launches}; T=${2:-r02}
BENCH="python bench.py --steps 2 --warmup 1 --no-cpu --no-dense"
case $MODE in
  launches)
    $BENCH > gpurun_out/${T}_plain.log 2>&1 &&
    ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/${T}_launches.csv $BENCH > gpurun_out/${T}_ncu.log 2>&1 ;;
  frame)
    BENCH="python bench.py --steps 1 --warmup 1 --no-cpu --no-dense"
    B200RT_NO_GRAPH=1 $BENCH > gpurun_out/${T}_plain.log 2>&1 &&
    B200RT_NO_GRAPH=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio --clock-control none -c 160 --csv --log-file gpurun_out/${T}_frame.csv $BENCH > gpurun_out/${T}_ncu.log 2>&1 ;;
  dense)
    BENCH="python tools/one_frame.py dense"
    $BENCH > gpurun_out/${T}_plain.log 2>&1 &&
    ncu --set full --clock-control none --import-source on -k regex:"k_isect3" -s 16 -c 8 -o gpurun_out/${T}_dense $BENCH > gpurun_out/${T}_ncu.log 2>&1 ;;
  culled)
    BENCH="python tools/one_frame.py culled"
    B200RT_NO_GRAPH=1 $BENCH > gpurun_out/${T}_plain.log 2>&1 &&
    B200RT_NO_GRAPH=1 ncu --set full --clock-control none --import-source on -k regex:"k_shadow_shade|k_closest_grid|k_setup|k_isect3|k_key|k_raygen" -s 9 -c 10 -o gpurun_out/${T}_culled $BENCH > gpurun_out/${T}_ncu.log 2>&1 ;;
esac
echo "rc=$?"; tail -3 gpurun_out/${T}_ncu.log | cut -c1-200
