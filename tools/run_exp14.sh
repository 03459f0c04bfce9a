for c in 8 7 6 4; do
  export YV3_EXTRA_NVCC="-DYV3_NMS_CLUSTER=$c"
  timeout 600 python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
  echo "cluster $c"
  timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "nms or fused_filter" 2>&1 | tail -1
  timeout 200 python bench.py --no-cpu-baseline --steps 20 --workload cfg5_512_c0_b16 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('cfg5', d['ms_per_step'], d['value'], d['stage_ms']['filter_plus_nms'])"
done
