cd $GRAFT_REPO_ROOT
for x in 15,3,2,6 12,3,2,6 16,4,2,6 15,3,2,4 15,3,3,6 15,3,2,8 12,4,2,6 15,5,2,6; do
  for m in coresident coresident_hist_main; do
    timeout 200 python bench.py --steps 50 --warmup 4 --no-aux --no-cpu-baseline --no-parity --overlap $m --coresident $x 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$m $x', round(d['ms_per_step'],5), round(d['value']/1e9,2))"
  done
done
