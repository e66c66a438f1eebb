cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python tools/quick_bench.py "$@" > gpurun_out/quick.log 2>&1
B200LSAP_LIB=$PWD/build_timing/var1t.so timeout 600 python tools/quick_bench.py c4t > gpurun_out/quick_timing.log 2>&1
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/quick.log gpurun_out/quick_timing.log
