# developer experiment: stream join once per graph vs every step
for args in "--pipeline 0 --graph-steps 1" "--pipeline 1 --graph-steps 4" "--pipeline 1 --graph-steps 8" "--pipeline 1 --graph-steps 8 --force-graph" "--pipeline 1 --graph-steps 25 --force-graph" "--pipeline 1 --no-graph"; do
  timeout 200 python bench.py --no-cpu-baseline --steps 200 $args 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$args', '|', d['ms_per_step'], d['value'], d['config']['launch'], d['config']['stream_join'], d['e2e']['ms_per_step'])"
done
