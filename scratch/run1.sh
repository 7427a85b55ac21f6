set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu2.log
python benchmarks/bench_quant.py --quick > gpurun_out/bq_mode1.log 2>&1
