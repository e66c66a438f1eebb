cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
for sc in 0.125 0.25; do echo "== eps scale $sc"; B200LSAP_EPS_SCALE=$sc timeout 600 python tools/quick_bench.py seeds 2>&1 | grep -v "full bid"; done > gpurun_out/sweep_eps.log 2>&1
cat gpurun_out/sweep_eps.log
