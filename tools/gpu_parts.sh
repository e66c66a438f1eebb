cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for p in 16 8 4 1; do echo "== max_parts $p"; B200LSAP_MAX_PARTS=$p timeout 300 python tools/quick_bench.py c4d sweep; done > gpurun_out/quick.log 2>&1
cat gpurun_out/quick.log
