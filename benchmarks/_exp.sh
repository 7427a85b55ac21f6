timeout 600 python -m pytest tests/test_gemm_w8a16_gpu.py -m gpu -x -q 2>&1 | tail -2
for t in 0 1; do echo "== VXK_W16_TMEM_A=$t"; VXK_W16_TMEM_A=$t python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,128; done
