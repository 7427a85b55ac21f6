cd $GRAFT_REPO_ROOT
N=$1
timeout 300 python -m pytest tests/test_multigpu_gpu.py -x -q > gpurun_out/r27_mgpu_test.log 2>&1; tail -30 gpurun_out/r27_mgpu_test.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r01e_bench_n$N.json 2> gpurun_out/r27_bench_n$N.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r01e_bench_n$N.json").read().strip().splitlines()[-1])
print(d["value"], d["gemm"]["reassembled"])
PY
tail -3 gpurun_out/r27_bench_n$N.err
