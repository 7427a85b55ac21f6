set -x
timeout 300 python benchmarks/gemm_probe.py 8192 8192 8192 > gpurun_out/probe1.log 2>&1
timeout 300 python benchmarks/gemm_probe.py 8192 8192 28672 --kinds bf16_cublas,fp8_cublaslt,fp8_w8a8,mxfp4 >> gpurun_out/probe1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -c 2 -o gpurun_out/prof_gemm_pair_r01 -f python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.05 --kinds fp8_w8a8,mxfp4 > gpurun_out/ncu_pair.log 2>&1
