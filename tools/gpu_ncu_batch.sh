cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python tools/quick_bench.py c3s > gpurun_out/c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_auction_batch -s 1 -c 1 -o gpurun_out/r1_batch python tools/quick_bench.py c3s > gpurun_out/ncu_batch.log 2>&1
cat gpurun_out/c3_plain.log; tail -3 gpurun_out/ncu_batch.log
