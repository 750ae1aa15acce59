# per-rank time of an emulated 8-GPU frame for several strip widths: max over ranks = the frame time
cd $GRAFT_REPO_ROOT
for sw in 16 480 240 64; do
  for r in 0 1 2 3 4 5 6 7; do
    B200RT_STRIP=$sw B200RT_EMULATE_WORLD=8 B200RT_EMULATE_RANK=$r timeout 120 python bench.py --steps 5 --warmup 2 --no-cpu --no-dense 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('strip $sw rank $r ms', round(d['ms_per_step'],3))"
  done
done
