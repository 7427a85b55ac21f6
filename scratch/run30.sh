cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gemm_w8a16_gpu.py -x -q -k small_m 2>&1 | tail -2
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,64,128 > gpurun_out/r30_sweep_w8a16.log 2>&1; tail -12 gpurun_out/r30_sweep_w8a16.log
