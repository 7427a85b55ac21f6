timeout 900 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -2
timeout 300 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_fp8 --Ms 1,64,128,256 2>&1 | tail -16
echo PDL=0
VXK_SKINNY_PDL=0 timeout 300 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_fp8 --Ms 1,64 2>&1 | tail -8
echo EAGER
timeout 300 python benchmarks/bench_gemm_sweep.py --sets llama8b_fp8 --Ms 1,64 2>&1 | tail -8
