cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 900 python tools/quick_bench.py struct c4t sq
} > gpurun_out/struct.log 2>&1
cat gpurun_out/struct.log
