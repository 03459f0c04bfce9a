for c in 1 2 3 6; do for d in 0 64; do echo "cpw $c debug $d"; YV3_CPW=$c YV3_DEBUG=$d timeout 120 python tools/dense_bench.py cfg2_416_c80_b64 60 2>&1 | cut -c1-110; done; done
