for v in 256 128; do echo "== VXK_SKINNY_MAX_M=$v"; VXK_SKINNY_MAX_M=$v python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_fp8 --Ms 160,192,256; done
