# round-end evidence: plain bench lines first (never under a profiler), then the ncu captures of the same code
set -x
timeout 600 python bench.py > gpurun_out/final_bench.log 2>&1; tail -1 gpurun_out/final_bench.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_ref.log 2>&1; tail -1 gpurun_out/final_ref.log | cut -c1-200
timeout 300 python bench.py --no-cpu-baseline --fused 0 > gpurun_out/final_separate.log 2>&1; tail -1 gpurun_out/final_separate.log | cut -c1-200
timeout 300 python bench.py --no-cpu-baseline --layout heads > gpurun_out/final_heads.log 2>&1; tail -1 gpurun_out/final_heads.log | cut -c1-200
export YV3_PROFILE_FUSED=1
timeout 120 python tools/profile_step.py && timeout 600 ncu --set full --clock-control none --import-source on --launch-skip 14 --launch-count 7 -o gpurun_out/prof_r01_final_step -f python tools/profile_step.py > gpurun_out/ncu_final_step.log 2>&1; tail -1 gpurun_out/ncu_final_step.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01_final.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_bench.log 2>&1; tail -1 gpurun_out/ncu_bench.log | cut -c1-100
