timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for e in 0 1; do
  if [ $e = 1 ]; then export YV3_NO_CLUSTER=1; fi
  echo "no_cluster=$e"
  timeout 200 python bench.py --no-cpu-baseline --steps 200 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg2', d['ms_per_step'], d['value'])"
  timeout 200 python bench.py --no-cpu-baseline --steps 20 --workload cfg5_512_c0_b16 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg5', d['ms_per_step'], d['value'], d['stage_ms']['filter_plus_nms'])"
  timeout 200 python bench.py --no-cpu-baseline --steps 20 --workload cfg4_416_c80_b256_dense 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg4', d['ms_per_step'], d['value'], d['stage_ms']['filter_plus_nms'])"
done
