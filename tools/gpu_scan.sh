# k_bid variants: libraries under build_timing/ x B200LSAP_SWEEP_SPLIT settings
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
for sp in 1 0; do
  echo "== base split=$sp"; B200LSAP_SWEEP_SPLIT=$sp timeout 300 python tools/quick_bench.py sweep2 2>&1 | grep sweep2
done
for f in "$@"; do
  for sp in 0; do
    echo "== $f split=$sp"; B200LSAP_SWEEP_SPLIT=$sp B200LSAP_LIB=$PWD/build_timing/$f.so timeout 300 python tools/quick_bench.py sweep2 2>&1 | grep sweep2
    B200LSAP_LIB=$PWD/build_timing/$f.so timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
  done
done
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python tools/quick_bench.py c1 c2 sq lsap 2>&1 | grep -v "full bid"
} > gpurun_out/scan.log 2>&1
cat gpurun_out/scan.log
