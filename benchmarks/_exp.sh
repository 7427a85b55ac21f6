for d in 0 4 6; do echo "== VXK_GG_DEBUG=$d"; VXK_GG_DEBUG=$d python benchmarks/bench_grouped_gemm.py --graph --Ms 256 --kinds fp8,mxfp4; done
python benchmarks/bench_grouped_gemm.py --graph --Ms 256 --kinds bf16
