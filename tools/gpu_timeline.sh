# per-launch timeline (B200RT_TIMELINE) of one frame: bash tools/gpu_timeline.sh [workload] [emulate_world]
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
B200RT_TIMELINE=1 B200RT_EMULATE_WORLD=${2:-0} timeout 300 python bench.py --workload ${1:-c4} --steps 1 --warmup 3 --no-cpu > gpurun_out/tl.json 2> gpurun_out/tl.err
# frames: stats, 3 warmup, 1 timed, stats, e2e...: print the 5th frame's block (timed, concurrent)
python - <<'PY'
lines=[l for l in open('gpurun_out/tl.err') if 'timeline' in l]
# split into frames at 'closest<1> d0' or first line pattern
frames=[]; cur=[]
for l in lines:
    if ('closest<1> d0' in l or 'closest<0> d0' in l) and cur: frames.append(cur); cur=[]
    cur.append(l)
if cur: frames.append(cur)
print(len(frames),'frames')
f=frames[4] if len(frames)>4 else frames[-1]
print(''.join(f))
PY
