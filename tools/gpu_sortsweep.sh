cd $GRAFT_REPO_ROOT
run() { B200RT_EMULATE_WORLD=$1 B200RT_SORT_POSBITS=$2 B200RT_SORT_DIRN=$3 timeout 300 python bench.py --workload c4 --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('emu $1 posbits $2 dirn $3 ms', round(d['ms_per_step'],3))"; }
for c in "8 10 2" "8 9 2" "8 11 2" "8 10 3" "8 10 1" "8 9 1" "4 11 2" "4 11 3" "4 12 2" "2 12 4" "2 12 3" "2 12 2" "2 13 3" "0 13 3" "0 13 2" "0 14 2" "0 14 3" "0 12 4"; do run $c; done
