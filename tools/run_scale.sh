# one multi-GPU record: gpurun --gpus N --timeout 900 -- 'bash tools/run_scale.sh N [tag]'
N=${1:-2}
TAG=${2:-r02}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi_rank.py -x -q -m gpu 2>&1 | tail -2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/${TAG}_bench_n$N.log 2>&1
echo "bench exit $?"; tail -1 gpurun_out/${TAG}_bench_n$N.log | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 200 --warmup 10 --workload cfg3_608_c80_b64 --no-extras > gpurun_out/${TAG}_bench_cfg3_n$N.log 2>&1
echo "cfg3 exit $?"; tail -1 gpurun_out/${TAG}_bench_cfg3_n$N.log | cut -c1-300
