cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== batch tests"; timeout 600 python -m pytest tests/test_gpu_batch_device.py -m gpu -x -q 2>&1 | tail -15
echo "== filter kernel"; timeout 300 python tools/quick_bench.py c3
echo "== streaming kernel (no filter)"; B200LSAP_BATCH_FILTER=0 timeout 300 python tools/quick_bench.py c3
} > gpurun_out/filter.log 2>&1
cat gpurun_out/filter.log
if [ "$1" = "ncu" ]; then
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_auction_batch -s 1 -c 1 -o gpurun_out/${2:-r1i}_batch python tools/quick_bench.py c3s > gpurun_out/ncu_batch.log 2>&1
tail -3 gpurun_out/ncu_batch.log
fi
