# developer experiment: CTA size of the dense pass
for t in 256 128 64; do
  export YV3_EXTRA_NVCC="-DYV3_DENSE_THREADS=$t"
  timeout 600 python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
  grep -A2 "loss_dense_kernelILi1ELi1ELi0" yolov3-tensorflow_b200/lib/build.log | grep -E "Used|spill" | head -2
  for c in 1 2 4; do
    echo "threads $t cpw $c"; YV3_CPW=$c timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 100 2>&1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['loss_dense'], d['loss_all']['median_us'], d['filter']['median_us'])"
  done
done
