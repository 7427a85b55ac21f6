python -m pytest tests/test_quant_parity_gpu.py -x -q 2>&1 | tail -4
python benchmarks/bench_quant.py --quick 2>&1
