cd $GRAFT_REPO_ROOT
( timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/r35_tests.log 2>&1
grep -E "passed|failed|error" gpurun_out/r35_tests.log | tail -3
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 300 python bench.py 2> gpurun_out/r35_bench.err | cut -c1-400
timeout 300 python bench.py --impl reference 2>> gpurun_out/r35_bench.err | cut -c1-200
