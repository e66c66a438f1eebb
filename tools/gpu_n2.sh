cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
for f in "$@"; do echo "== $f"; B200LSAP_LIB=$PWD/build_timing/$f.so timeout 300 python tools/quick_bench.py c4t sq 2>&1 | grep -v "full bid"; done
B200LSAP_LIB=$PWD/build_timing/n4.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
} > gpurun_out/n2.log 2>&1
cat gpurun_out/n2.log
