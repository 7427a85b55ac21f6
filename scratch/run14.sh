timeout 900 python -m pytest tests/test_quant_parity_gpu.py -x -q 2>&1 | tail -2
timeout 300 python benchmarks/bench_quant.py --T 8192 --only rms_mxfp4,rms_block_fp8_128 | grep -E "T= 8192"
