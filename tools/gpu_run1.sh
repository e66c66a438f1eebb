set -x
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python tools/quick_bench.py c1 c2 c3 sq > gpurun_out/quick.log 2>&1
timeout 300 python tools/quick_bench.py sweep > gpurun_out/sweep_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_bid -s 1 -c 2 -o gpurun_out/r1b_kbid python tools/quick_bench.py sweep > gpurun_out/ncu_kbid.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/quick.log
