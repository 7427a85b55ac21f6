cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r17_tests.log 2>&1
tail -3 gpurun_out/r17_tests.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/r17_smoke.log 2>&1; tail -2 gpurun_out/r17_smoke.log
( time timeout 600 python bench.py ) > gpurun_out/r17_bench.log 2>&1; tail -4 gpurun_out/r17_bench.log
( time timeout 600 python bench.py --impl reference ) > gpurun_out/r17_bench_ref.log 2>&1; tail -3 gpurun_out/r17_bench_ref.log
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_fp8 --Ms 1,16,64,128,256,1024,8192 > gpurun_out/r17_sweep8b.log 2>&1; tail -30 gpurun_out/r17_sweep8b.log
