# Round check on a B200 box: GPU tests, smoke, plain bench, reference arm, then the ncu launch list of the same bench command.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
TAG=${1:-r1}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_${TAG}.json 2> gpurun_out/bench_${TAG}.err; echo "bench rc=$?" >> gpurun_out/bench_${TAG}.err
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref_${TAG}.json 2> gpurun_out/bench_ref_${TAG}.err
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/bench_under_ncu.log 2>&1
tail -4 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cut -c1-2500 gpurun_out/bench_${TAG}.json; tail -3 gpurun_out/bench_${TAG}.err; cut -c1-600 gpurun_out/bench_ref_${TAG}.json
