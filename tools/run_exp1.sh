# developer experiment: stream priorities for the two halves of a step; 64-register build of the dense pass
set -x
for p in none post loss; do
  timeout 200 python bench.py --no-cpu-baseline --steps 200 --priority $p 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('PRIORITY', d['config']['high_priority_stream'], d['ms_per_step'], d['config']['launch'], d['stage_ms'])"
done
timeout 200 python bench.py --no-cpu-baseline --steps 200 --priority post --force-graph 2>&1 | tail -1 | cut -c1-400
timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 100 2>&1 | cut -c1-600
export YV3_EXTRA_NVCC="-DYV3_DENSE_MINB=4"
timeout 600 python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 100 2>&1 | cut -c1-600
timeout 200 python bench.py --no-cpu-baseline --steps 200 --priority post 2>&1 | tail -1 | cut -c1-300
