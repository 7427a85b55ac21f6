set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/bench4.json 2> gpurun_out/bench4.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"row_quant_kernel|flat_quant_kernel" -s 4 -c 2 -o gpurun_out/prof_quant_r01c -f python bench.py --steps 2 --warmup 1 > gpurun_out/ncu_q.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -c 2 -o gpurun_out/prof_gemm_r01c -f python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.02 --kinds fp8_w8a8,mxfp4 > gpurun_out/ncu_g.log 2>&1
ncu --set full --clock-control none -k regex:gemm_tc2 -c 1 -o gpurun_out/prof_gemm_blk_r01c -f python benchmarks/gemm_probe.py 4096 7168 16384 --secs 0.02 --kinds fp8_block_f32 > gpurun_out/ncu_gb.log 2>&1
