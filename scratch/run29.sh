cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r29_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r29_tests.log | tail -3
timeout 600 python benchmarks/bench_gemm_sweep.py --sets dsv3_blockfp8 --Ms 1024,8192 > gpurun_out/r29_sweep_blk.log 2>&1; tail -12 gpurun_out/r29_sweep_blk.log
