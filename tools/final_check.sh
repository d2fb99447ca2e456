mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 60 python tools/interaction_scatter_bench.py > gpurun_out/interaction_scatter.json 2> gpurun_out/interaction_scatter.err; cat gpurun_out/interaction_scatter.json
timeout 200 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"; python tools/print_bench.py gpurun_out/bench_n1.json
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
