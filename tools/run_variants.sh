# developer helper: schedule variants of the fused step (GPU box)
for v in "--fused-depth 2" "--fused-depth 3" "--fused-depth 4" "--priority none"; do
  timeout 200 python bench.py --no-cpu-baseline $v 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('$v', '%.2f us' % (d['ms_per_step']*1e3), d['config']['launch'][:40])"
done
