for args in "--serialize-sweeps 0" "--serialize-sweeps 1" "--serialize-sweeps 1 --priority none" "--serialize-sweeps 1 --graph-steps 25 --force-graph"; do
  timeout 200 python bench.py --no-cpu-baseline --steps 200 $args 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$args', '|', d['ms_per_step'], d['value'], d['config']['launch'])"
done
