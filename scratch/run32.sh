cd $GRAFT_REPO_ROOT
( timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r32_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r32_tests.log | tail -3
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,128,256,1024,4096,8192 > gpurun_out/r01e_sweep_w8a16.txt 2>&1; grep "M=    1 \|M=   64 " gpurun_out/r01e_sweep_w8a16.txt
timeout 300 python bench.py > gpurun_out/r01e_bench_n1.json 2> gpurun_out/r32_bench.err; tail -c 300 gpurun_out/r01e_bench_n1.json
