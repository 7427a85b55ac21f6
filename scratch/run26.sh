cd $GRAFT_REPO_ROOT
N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r01e_bench_n$N.json 2> gpurun_out/r26_bench_n$N.err
tail -c 1500 gpurun_out/r01e_bench_n$N.json
tail -3 gpurun_out/r26_bench_n$N.err
