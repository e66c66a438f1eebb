cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_sparse.py -x -q 2>&1 | tail -15
timeout 600 python tools/quick_bench.py sparse 2>&1 | tail -12
} > gpurun_out/sparse.log 2>&1
cat gpurun_out/sparse.log
