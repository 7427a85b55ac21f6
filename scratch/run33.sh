cd $GRAFT_REPO_ROOT
for g in 4 8 16 32; do echo GROUP_M=$g; VXK_GEMM_GROUP_M=$g timeout 300 python benchmarks/bench_gemm_sweep.py --sets llama70b_mxfp4,llama8b_fp8 --Ms 8192 2>&1 | tail -8; done > gpurun_out/r33_groupm.log 2>&1
echo NACC2; VXK_GEMM_NACC=2 timeout 300 python benchmarks/bench_gemm_sweep.py --sets llama70b_mxfp4 --Ms 8192 2>&1 | tail -4 >> gpurun_out/r33_groupm.log
cat gpurun_out/r33_groupm.log
