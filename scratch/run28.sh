cd $GRAFT_REPO_ROOT
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r28_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r28_tests.log | tail -3
timeout 300 python __graft_entry__.py smoke > gpurun_out/r28_smoke.log 2>&1; tail -1 gpurun_out/r28_smoke.log
timeout 600 python bench.py > gpurun_out/r01e_bench_n1.json 2> gpurun_out/r28_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/r01e_bench_ref.json 2>> gpurun_out/r28_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01e_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r28_ncu_bench.log 2>&1
python - <<PY
import json
d=json.loads(open("gpurun_out/r01e_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["roofline"]["frac"], d["e2e"]["value"])
for r in d["gemm"]["rows"]: print(r["kind"], r["M"],r["N"],r["K"], round(r["tflops"]), round(r["frac_of_dense_peak"],3))
PY
