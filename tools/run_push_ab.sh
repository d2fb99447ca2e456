mkdir -p gpurun_out
timeout 250 python -m pytest tests/test_gpu_modules.py -m gpu -q -x -k "multi_gpu" > gpurun_out/t_push.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/t_push.log
for P in 1 0; do
  B200_PEER_PUSH=$P timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2970$P bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n2_push$P.json 2> gpurun_out/bench_n2_push$P.err; echo "push=$P rc=$?"
  python tools/print_bench.py gpurun_out/bench_n2_push$P.json
done
