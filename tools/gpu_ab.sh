# A/B of a side build (build_timing/$1.so) against the tree's build: batch parity tests and the C3 timing
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== $1: batch tests"; B200LSAP_LIB=$PWD/build_timing/$1.so timeout 600 python -m pytest tests/test_gpu_batch_device.py -m gpu -x -q 2>&1 | tail -4
echo "== $1: c3"; B200LSAP_LIB=$PWD/build_timing/$1.so timeout 300 python tools/quick_bench.py c3
echo "== tree build: c3"; timeout 300 python tools/quick_bench.py c3
echo "== $1: c3 again"; B200LSAP_LIB=$PWD/build_timing/$1.so timeout 300 python tools/quick_bench.py c3
} > gpurun_out/ab_$1.log 2>&1
cat gpurun_out/ab_$1.log
