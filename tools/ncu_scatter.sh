mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_tbe.py -m gpu -q -k "interaction_backward_scatter" > gpurun_out/t_scatter.log 2>&1; echo "scatter tests rc=$?"; tail -2 gpurun_out/t_scatter.log
timeout 240 ncu --set full --clock-control none --import-source on -k regex:interaction_backward_mma --launch-skip 6 -c 2 -o gpurun_out/r1_scatter python tools/interaction_scatter_bench.py > gpurun_out/ncu_scatter.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_scatter.log; ls -la gpurun_out/*.ncu-rep
