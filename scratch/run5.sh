timeout 300 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -3
VXK_GEMM_NACC=1 timeout 300 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -2
VXK_GEMM_NACC=2 timeout 300 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -2
P="timeout 120 python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.3 --kinds fp8_w8a8,mxfp8_block,mxfp4"
for n in 1 2; do echo "NACC=$n"; VXK_GEMM_NACC=$n $P; done
echo "NACC=2 no cp"; VXK_GEMM_NACC=2 VXK_GEMM_DEBUG=1 $P
echo "auto K=28672"; timeout 120 python benchmarks/gemm_probe.py 8192 8192 28672 --secs 0.3 --kinds fp8_w8a8,mxfp4
echo "auto N=57344"; timeout 120 python benchmarks/gemm_probe.py 8192 57344 8192 --secs 0.3 --kinds fp8_w8a8,mxfp4
