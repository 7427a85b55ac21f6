cd $GRAFT_REPO_ROOT
timeout 300 python -m pytest tests/test_gemm_w8a16_gpu.py -x -q -k small_m 2>&1 | tail -2
for pf in 0 6 12 24; do echo PF=$pf; VXK_W16_PREFETCH=$pf timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,64 2>&1 | tail -8; done > gpurun_out/r31_pf.log 2>&1
cat gpurun_out/r31_pf.log
