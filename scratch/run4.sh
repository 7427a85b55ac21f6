set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -c 1 -o gpurun_out/prof_mxfp4_pair_r01 -f python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.05 --kinds mxfp4 > gpurun_out/ncu_pair4.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:gemm_tc2 -c 1 -o gpurun_out/prof_mxfp8_pair_r01 -f python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.05 --kinds mxfp8_block > gpurun_out/ncu_pair8.log 2>&1
