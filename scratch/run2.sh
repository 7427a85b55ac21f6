set -x
timeout 600 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -25 > gpurun_out/pytest_gemm2.log
timeout 300 python benchmarks/bench_gemm_quick.py > gpurun_out/gemm_pair.log 2>&1
VXK_GEMM_SINGLE_CTA=1 timeout 300 python benchmarks/bench_gemm_quick.py > gpurun_out/gemm_single.log 2>&1
