# fused-sweep evidence: parity tests, then plain bench lines (fused and separate), then the ncu captures
set -x
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/gpu_tests_fused.log; tail -3 gpurun_out/gpu_tests_fused.log
timeout 300 python bench.py --fused 1 > gpurun_out/fused_bench.log 2>&1; tail -1 gpurun_out/fused_bench.log | cut -c1-250
timeout 300 python bench.py --fused 0 --no-cpu-baseline > gpurun_out/sep_bench.log 2>&1; tail -1 gpurun_out/sep_bench.log | cut -c1-250
YV3_PROFILE_FUSED=1 timeout 120 python tools/profile_step.py && YV3_PROFILE_FUSED=1 timeout 600 ncu --set full --clock-control none --import-source on --launch-skip 14 --launch-count 7 -o gpurun_out/prof_r01_fused_step -f python tools/profile_step.py > gpurun_out/ncu_fused_step.log 2>&1; tail -1 gpurun_out/ncu_fused_step.log
