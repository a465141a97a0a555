# r02 call S (8 GPUs): one handle over eight GPUs (tests), torchrun bench as the driver launches it, H2D ceiling probe
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
nproc
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -3
timeout 480 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err; echo "exit $?"; grep -v "^$\|OMP_NUM\|^\*\*\*" gpurun_out/bench_8gpu.err | tail -3 | cut -c1-300; python profiles/r02/summ.py gpurun_out/bench_8gpu.json > gpurun_out/bench_8gpu.txt 2>&1; head -16 gpurun_out/bench_8gpu.txt
echo "== H2D probe, one process, all GPUs"; timeout 200 python profiles/h2d_probe.py 2>&1 | tail -6 | cut -c1-400
