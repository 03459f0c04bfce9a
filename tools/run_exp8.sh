for v in "256 3" "256 4" "128 6" "128 8" "128 7"; do
  set -- $v
  export YV3_EXTRA_NVCC="-DYV3_DENSE_THREADS=$1 -DYV3_DENSE_MINB=$2"
  timeout 600 python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
  grep -A2 "loss_dense_kernelILi1ELi1ELi0" yolov3-tensorflow_b200/lib/build.log | grep -E "Used|spill" | head -2
  echo "threads $1 minb $2"; timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 100 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['loss_dense'], d['loss_all']['median_us'], d['loss_fwd_only']['median_us'])"
done
