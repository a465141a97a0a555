# r02 call P (2 GPUs): final check of the multi-rank path: multi-GPU tests, torchrun bench as the driver launches it, bounded
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "exit $?"; grep -v "^$\|OMP_NUM\|^\*\*\*" gpurun_out/bench_2gpu.err | tail -3 | cut -c1-300; python profiles/r02/summ.py gpurun_out/bench_2gpu.json > gpurun_out/bench_2gpu.txt 2>&1; head -5 gpurun_out/bench_2gpu.txt; grep one_process gpurun_out/bench_2gpu.txt | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | cut -c1-300
