timeout 300 python -m pytest tests/test_gemm_parity_gpu.py -x -q 2>&1 | tail -3
P="timeout 120 python benchmarks/gemm_probe.py"
for m in 1 16 128; do $P $m 4096 4096 --secs 0.2 --kinds fp8_cublaslt,fp8_w8a8,mxfp4; $P $m 28672 4096 --secs 0.2 --kinds fp8_cublaslt,fp8_w8a8,mxfp4; $P $m 4096 14336 --secs 0.2 --kinds fp8_cublaslt,fp8_w8a8,mxfp4; done
for m in 256 512 1024 2048; do $P $m 6144 4096 --secs 0.2 --kinds fp8_cublaslt,fp8_w8a8,mxfp4; done
