cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== stream kernel, 2 lanes per problem"; B200LSAP_BATCH_LANES=2 timeout 300 python tools/quick_bench.py c3
} > gpurun_out/batch.log 2>&1
cat gpurun_out/batch.log
