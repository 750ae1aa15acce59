#!/bin/bash
# Final numbers of a build on one GPU: tests, smoke, bench lines of all configs + the reference arm, ncu passes, checked build.  usage: tools/gpu_final.sh [TAG]
cd "$(dirname "$0")/.."; mkdir -p gpurun_out; T=${1:-r02}
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/${T}_pytest_gpu_final.log 2>&1; tail -1 gpurun_out/${T}_pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; tail -1 gpurun_out/${T}_smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/${T}_bench_c4.json 2> gpurun_out/${T}_bench_c4.err
for wl in c1 c2 c3; do timeout 300 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu > gpurun_out/${T}_bench_$wl.json 2> gpurun_out/${T}_bench_$wl.err; done
timeout 300 python bench.py --workload c5 --steps 64 --warmup 3 --no-cpu > gpurun_out/${T}_bench_c5.json 2> gpurun_out/${T}_bench_c5.err
for wl in c4_s2048 c4_s4096; do timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu --no-dense > gpurun_out/${T}_bench_$wl.json 2> gpurun_out/${T}_bench_$wl.err; done
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_reference_arm.json 2> gpurun_out/${T}_bench_reference_arm.err
tools/gpu_profile.sh launches $T; tools/gpu_profile.sh frame $T; tools/gpu_profile.sh culled $T
tools/gpu_check.sh > /dev/null 2>&1; tail -4 gpurun_out/r02_check.log
python - $T <<'PY'
import json,sys
T=sys.argv[1]
for n in ["c4","c1","c2","c3","c5","c4_s2048","c4_s4096","reference_arm"]:
    try:
        j=json.loads(open(f"gpurun_out/{T}_bench_{n}.json").read().strip().splitlines()[-1])
        print(n, "ms", round(j["ms_per_step"],4), "value", round(j["value"],1), "e2e ms", round(j["e2e"].get("ms_per_step",0),4), "frac", (j.get("roofline") or {}).get("frac"), "launches", j.get("gpu_launches"))
    except Exception as e: print(n, "ERR", e)
PY
