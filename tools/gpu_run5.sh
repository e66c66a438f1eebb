cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -12
timeout 900 python tools/quick_bench.py "$@"
} > gpurun_out/run5.log 2>&1
cat gpurun_out/run5.log
