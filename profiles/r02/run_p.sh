# r02 call P (2 GPUs): host time per file submit under torchrun
export ICB_POOL_CACHE=/tmp/icb_pool
mkdir -p gpurun_out
ICB_DEC_DEBUG=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 4 --warmup 3 --no-legs --no-cpu --no-check > gpurun_out/bench_2gpu_dbg.json 2> gpurun_out/bench_2gpu_dbg.err; echo "exit $?"
grep "submit_jpg_file_fit" gpurun_out/bench_2gpu_dbg.err | tail -4; grep "icb decode" gpurun_out/bench_2gpu_dbg.err | sort | uniq -c | sort -rn | head -4
python profiles/r02/summ.py gpurun_out/bench_2gpu_dbg.json 2>/dev/null | head -3
echo "== N=1 launched through torchrun too"
ICB_DEC_DEBUG=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 1 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 1 --steps 4 --warmup 3 --no-legs --no-cpu --no-check > gpurun_out/bench_1gpu_dbg.json 2> gpurun_out/bench_1gpu_dbg.err; echo "exit $?"
grep "submit_jpg_file_fit" gpurun_out/bench_1gpu_dbg.err | tail -2
python profiles/r02/summ.py gpurun_out/bench_1gpu_dbg.json 2>/dev/null | head -3
