#!/bin/bash
# per-launch timeline of rank 0's share of an 8-GPU C4 frame (and of the whole frame), plain launches
cd "$(dirname "$0")/.."
B200RT_TIMELINE=1 B200RT_EMULATE_WORLD=8 python bench.py --no-cpu --no-dense --steps 1 --warmup 3 2> gpurun_out/tl8.err > /dev/null
grep timeline gpurun_out/tl8.err | tail -45 > gpurun_out/r02_timeline_emu8.txt
B200RT_TIMELINE=1 python bench.py --no-cpu --no-dense --steps 1 --warmup 3 2> gpurun_out/tl1.err > /dev/null
grep timeline gpurun_out/tl1.err | tail -45 > gpurun_out/r02_timeline_n1.txt
