#!/bin/bash
# DRAM traffic of the HBM-bound kernels, one `ncu --set full --clock-control none` capture each, run on the GPU box:
#     gpurun --timeout 900 -- 'bash tools/traffic_capture.sh'
# Every command first runs plain (must exit 0), then once more under ncu with only the launch after the warm-up captured.
# The reports land in gpurun_out/traffic_*.ncu-rep; tools/traffic_json.py turns them into profiles/traffic.json (read by
# bench.py for roofline.traffic) and tools/ncu_summary.py prints the per-kernel summaries kept under profiles/.
set -u
mkdir -p gpurun_out
NCU="ncu --set full --clock-control none --import-source on"

capture() {     # name, kernel regex, probe arguments...
    local name=$1 kern=$2
    shift 2
    if ! timeout 300 python tools/sweep_probe.py "$@" --once > gpurun_out/traffic_${name}.plain.log 2>&1; then
        echo "plain run of $name failed"; tail -5 gpurun_out/traffic_${name}.plain.log; return 1
    fi
    # 3 warm-up launches, the 4th is the one captured
    timeout 600 $NCU -k regex:$kern --launch-skip 3 --launch-count 1 -f -o gpurun_out/traffic_${name} \
        python tools/sweep_probe.py "$@" --once > gpurun_out/traffic_${name}.ncu.log 2>&1
    echo "$name: exit $?"
    tail -n 2 gpurun_out/traffic_${name}.ncu.log
}

capture cfg2_fused  fused_sweep_kernel --size 416 --classes 80 --batch 64  --kind fused  --thresh 0.1  --focal 0 --rect 0
capture cfg2_dense  loss_dense_kernel  --size 416 --classes 80 --batch 64  --kind dense  --focal 0 --rect 0
capture cfg2_decode ^decode_kernel     --size 416 --classes 80 --batch 64  --kind decode --focal 0 --rect 0
capture cfg3_fused  fused_sweep_kernel --size 608 --classes 80 --batch 64  --kind fused  --thresh 0.1  --focal 0 --rect 0
capture cfg5_fused  fused_sweep_kernel --size 512 --classes 0  --batch 128 --kind fused  --thresh 0.95 --focal 1 --rect 1
ls -la gpurun_out/traffic_*.ncu-rep
