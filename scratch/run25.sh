cd $GRAFT_REPO_ROOT
( time timeout 900 python -m pytest tests/test_gemm_w8a16_gpu.py -x -q ) > gpurun_out/r25_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r25_tests.log | tail -3
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,128,256 > gpurun_out/r25_sweep_w8a16.log 2>&1; tail -24 gpurun_out/r25_sweep_w8a16.log
