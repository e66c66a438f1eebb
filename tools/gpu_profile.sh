# ncu --set full captures of the three kernels that matter (each after its plain run exited 0); reports come
# back in gpurun_out/ and are summarised into profiles/ by tools/summarize_ncu.py
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TAG=${1:-r1e}
timeout 300 python tools/quick_bench.py sweep2 > gpurun_out/sweep_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_bid -s 12 -c 1 -f -o gpurun_out/${TAG}_kbid python tools/quick_bench.py sweep2 > gpurun_out/ncu_kbid.log 2>&1
timeout 300 python tools/quick_bench.py c3s > gpurun_out/c3_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_auction_batch -s 1 -c 1 -f -o gpurun_out/${TAG}_batch python tools/quick_bench.py c3s > gpurun_out/ncu_batch.log 2>&1
timeout 300 python tools/quick_bench.py c4t > gpurun_out/c4_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_narrow -s 3 -c 1 -f -o gpurun_out/${TAG}_knarrow python tools/quick_bench.py c4t > gpurun_out/ncu_knarrow.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_auction_persistent -s 3 -c 1 -f -o gpurun_out/${TAG}_kpersistent python tools/quick_bench.py c4t > gpurun_out/ncu_kpersistent.log 2>&1
timeout 300 python tools/quick_bench.py sparse20k > gpurun_out/sparse_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_auction_csr -s 1 -c 1 -f -o gpurun_out/${TAG}_kcsr python tools/quick_bench.py sparse20k > gpurun_out/ncu_kcsr.log 2>&1
cat gpurun_out/sweep_plain.log gpurun_out/c3_plain.log gpurun_out/c4_plain.log gpurun_out/sparse_plain.log; ls -la gpurun_out/*.ncu-rep
