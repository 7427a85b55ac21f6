cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r24_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r24_tests.log | tail -3
timeout 300 python __graft_entry__.py smoke > gpurun_out/r24_smoke.log 2>&1; tail -1 gpurun_out/r24_smoke.log
timeout 600 python bench.py > gpurun_out/r01e_bench_n1.json 2> gpurun_out/r24_bench.err; tail -c 600 gpurun_out/r01e_bench_n1.json
timeout 600 python bench.py --impl reference > gpurun_out/r01e_bench_ref.json 2>> gpurun_out/r24_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01e_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r24_ncu_bench.log 2>&1
timeout 600 python benchmarks/bench_gemm_sweep.py --graph --sets llama8b_w8a16 --Ms 1,16,64,128,256,1024,4096,8192 > gpurun_out/r01e_sweep_w8a16.txt 2>&1; grep "M=  256" gpurun_out/r01e_sweep_w8a16.txt
timeout 900 python benchmarks/bench_gemm_sweep.py --graph --Ms 1,16,64,256,1024,8192 > gpurun_out/r01e_sweep_gemm.txt 2>&1; tail -5 gpurun_out/r01e_sweep_gemm.txt
timeout 600 python benchmarks/bench_quant.py > gpurun_out/r01e_bench_quant.txt 2>&1; grep "T= 8192" gpurun_out/r01e_bench_quant.txt | head -40
for shp in "16 28672 4096 w16sk" "8192 6144 4096 w16big"; do set -- $shp
  ncu --set full --clock-control none --import-source on -k regex:gemm_w8a16 -s 3 -c 1 -o /tmp/$4 -f python scratch/prof_w16.py $1 $2 $3 > gpurun_out/r24_ncu_$4.log 2>&1
  python tools/ncu_summary.py /tmp/$4.ncu-rep > gpurun_out/r01e_w8a16_$4.csv
done
