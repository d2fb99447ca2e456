mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_modules.py -m gpu -q -x -k "multi_gpu and peer-4" > gpurun_out/t_n4.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_n4.log
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29704 bench.py --gpus 4 --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "bench rc=$?"
python tools/print_bench.py gpurun_out/bench_n4.json
