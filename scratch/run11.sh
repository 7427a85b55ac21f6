P="timeout 60 python benchmarks/gemm_probe.py 8192 8192 8192 --secs 0.3 --kinds fp8_block_f32"
for d in 0 16 64 80; do echo "debug=$d"; VXK_GEMM_DEBUG=$d $P; done
