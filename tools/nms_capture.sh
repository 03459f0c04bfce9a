#!/bin/bash
# ncu --set full captures (with source) of the post-process kernels at the two NMS-bound workloads, on the GPU box:
#     gpurun --timeout 900 -- 'bash tools/nms_capture.sh'
# The NMS kernels are launched from the host for these runs (YV3_NMS_SKIP=8): ncu does not see device-launched grids.
set -u
mkdir -p gpurun_out
export YV3_PROFILE_FUSED=1 YV3_NMS_SKIP=8
NCU="ncu --set full --clock-control none --import-source on"
cap() {   # name, workload, kernel regex
    timeout 300 python tools/profile_step.py $2 3 > gpurun_out/nmscap_$1.plain.log 2>&1 || { echo "plain $1 failed"; tail -3 gpurun_out/nmscap_$1.plain.log; return 1; }
    timeout 600 $NCU -k regex:$3 --launch-skip 2 --launch-count 1 -f -o gpurun_out/nmscap_$1 python tools/profile_step.py $2 3 > gpurun_out/nmscap_$1.ncu.log 2>&1
    echo "$1: exit $?"; tail -n 1 gpurun_out/nmscap_$1.ncu.log
}
cap cfg4_sort    cfg4_416_c80_b256_dense sort_kernel
cap cfg4_gather  cfg4_416_c80_b256_dense gather_kernel
cap cfg4_warp    cfg4_416_c80_b256_dense nms_warp_kernel
cap stress_cluster cfg5_nms_stress_b16   nms_cluster_kernel
cap stress_sort  cfg5_nms_stress_b16     sort_kernel
