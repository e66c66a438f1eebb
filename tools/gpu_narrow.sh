# k_narrow experiments: plain library, then timing builds under build_timing/
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== base"; timeout 300 python tools/quick_bench.py c4t sq 2>&1 | grep -v "full bid"
for f in "$@"; do echo "== $f"; B200LSAP_LIB=$PWD/build_timing/$f.so timeout 300 python tools/quick_bench.py c4t 2>&1 | grep -v "full bid"; done

} > gpurun_out/narrow.log 2>&1
cat gpurun_out/narrow.log
