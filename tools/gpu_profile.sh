cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 900 $CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_r1.csv $CMD > gpurun_out/bench_under_ncu.log 2>&1
timeout 300 python tools/quick_bench.py sweep > gpurun_out/sweep_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_bid -s 1 -c 2 -o gpurun_out/r1_kbid python tools/quick_bench.py sweep > gpurun_out/ncu_kbid.log 2>&1
timeout 300 python tools/quick_bench.py c3s > gpurun_out/c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_auction_batch -s 1 -c 1 -o gpurun_out/r1_batch python tools/quick_bench.py c3s > gpurun_out/ncu_batch.log 2>&1
cat gpurun_out/bench_plain.json | cut -c1-600; wc -l gpurun_out/launches_r1.csv; cat gpurun_out/sweep_plain.log gpurun_out/c3_plain.log
