cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
{
echo "== batch tests"; timeout 600 python -m pytest tests/test_gpu_batch_device.py -m gpu -x -q 2>&1 | tail -15
echo "== filter kernel"; timeout 300 python tools/quick_bench.py c3
echo "== streaming kernel (no filter)"; B200LSAP_BATCH_FILTER=0 timeout 300 python tools/quick_bench.py c3
echo "== sanitizer, batch case"; timeout 600 compute-sanitizer --tool memcheck python tools/sanitize_cases.py batch 2>&1 | tail -8
echo "== sanitizer racecheck, batch case"; timeout 600 compute-sanitizer --tool racecheck python tools/sanitize_cases.py batch 2>&1 | tail -8
} > gpurun_out/filter.log 2>&1
cat gpurun_out/filter.log
