# round check of the tree's build, then A/B of a side build (build_timing/$2.so) on the batch tests and the C3 timing
cd $GRAFT_REPO_ROOT
bash tools/gpu_full.sh $1 > gpurun_out/full_$1.log 2>&1
bash tools/gpu_filter.sh ncu $1 > /dev/null 2>&1; cp gpurun_out/filter.log gpurun_out/filter_$1.log
if [ -n "$2" ]; then
{
echo "== $2: batch tests"; B200LSAP_LIB=$PWD/build_timing/$2.so timeout 600 python -m pytest tests/test_gpu_batch_device.py -m gpu -x -q 2>&1 | tail -4
echo "== $2: c3"; B200LSAP_LIB=$PWD/build_timing/$2.so timeout 300 python tools/quick_bench.py c3
echo "== tree build: c3"; timeout 300 python tools/quick_bench.py c3
} > gpurun_out/ab_$2.log 2>&1
fi
tail -12 gpurun_out/full_$1.log | cut -c1-300; tail -8 gpurun_out/filter_$1.log; cat gpurun_out/ab_$2.log
