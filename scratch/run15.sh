for op in rms_token_fp8 rms_block_fp8_128 rms_mxfp4 group_fp8_128; do
ncu --set full --clock-control none -k regex:"row_quant_kernel|flat_quant_kernel" -s 6 -c 1 -o /tmp/prof_${op} -f python benchmarks/bench_quant.py --quick --only $op > gpurun_out/ncu_rowq.log 2>&1
python tools/ncu_summary.py /tmp/prof_${op}.ncu-rep > gpurun_out/r01d_${op}.csv
ncu -i /tmp/prof_${op}.ncu-rep --page details > gpurun_out/r01d_${op}_details.txt
done
ls -la gpurun_out
