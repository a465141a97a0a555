# r02 call S2 (8 GPUs): torchrun bench as the driver launches it, final code, bounded
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; echo "exit $?"; grep -v "^$\|OMP_NUM\|^\*\*\*" gpurun_out/bench_8gpu.err | tail -3 | cut -c1-300; python profiles/r02/summ.py gpurun_out/bench_8gpu.json > gpurun_out/bench_8gpu.txt 2>&1; head -4 gpurun_out/bench_8gpu.txt | cut -c1-300; grep one_process gpurun_out/bench_8gpu.txt | cut -c1-330
