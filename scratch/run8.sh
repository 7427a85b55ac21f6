set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 400 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_s2.json 2> gpurun_out/bench_s2.err
timeout 300 python benchmarks/bench_quant.py > gpurun_out/bench_quant_s2.txt 2>&1
tail -20 gpurun_out/bench_quant_s2.txt
