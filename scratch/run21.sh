cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_gemm_w8a16_gpu.py -x -q ) > gpurun_out/r21_tests.log 2>&1
tail -4 gpurun_out/r21_tests.log
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,128,1024,8192 > gpurun_out/r21_sweep_w8a16.log 2>&1; tail -30 gpurun_out/r21_sweep_w8a16.log
