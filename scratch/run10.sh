set -x
timeout 600 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -3
P="timeout 60 python benchmarks/gemm_probe.py"
$P 8192 8192 8192 --secs 0.3 --kinds fp8_block_f32,mxfp8_block
$P 4096 7168 16384 --secs 0.3 --kinds fp8_block_f32,mxfp8_block
$P 4096 24576 1536 --secs 0.3 --kinds fp8_block_f32,mxfp8_block
$P 8192 8192 8192 --secs 1.5 --kinds fp8_block_f32
