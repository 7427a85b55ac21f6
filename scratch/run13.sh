P="timeout 60 python benchmarks/gemm_probe.py"
for d in 0 1024; do for pp in 0 1; do echo "debug=$d PIPE=$pp";
VXK_GEMM_DEBUG=$d VXK_BLK_PIPE=$pp $P 8192 8192 8192 --secs 0.3 --kinds fp8_block_f32
done; done
