python benchmarks/bench_grouped_gemm.py --graph --Ms 16,256,1024,8192 --kinds fp8,mxfp4
python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,8192
