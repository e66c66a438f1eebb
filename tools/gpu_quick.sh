cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 600 python tools/quick_bench.py "$@" > gpurun_out/quick.log 2>&1
cat gpurun_out/quick.log
