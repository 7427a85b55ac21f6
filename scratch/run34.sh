cd $GRAFT_REPO_ROOT
timeout 600 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -2
timeout 300 python benchmarks/bench_gemm_sweep.py --sets llama70b_mxfp4 --Ms 1024,8192 2>&1 | tail -8
timeout 300 python bench.py > gpurun_out/r01e_bench_n1.json 2> gpurun_out/r34_bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r01e_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["roofline"]["frac"])
for r in d["gemm"]["rows"]: print(r["kind"], r["M"],r["N"],r["K"], round(r["tflops"]), round(r["frac_of_dense_peak"],3))
PY
