#!/bin/bash
# The one GPU-side runner (replaces the round-1 scratch launchers).  Run through gpurun from the repo root:
#     gpurun --timeout 1500 -- bash tools/gpu.sh <command> [args]
# Everything a command prints also lands in gpurun_out/<command>[_tag].log, which gpurun merges back.
#
#   quick  ARGS...         tools/quick_bench.py ARGS (scratch timings: c1 c2 c3 c3shards c4t sq sparse ...)
#   tests  [PYTEST ARGS]   pytest -m gpu (default: the whole GPU suite)
#   batch                  batch parity tests + the 8-shard C3 timing
#   ab     NAME ARGS...    quick_bench ARGS with build_timing/NAME.so, with the tree's build, and with NAME again
#   check  TAG             round check: GPU tests, smoke, bench.py, bench.py --impl reference, then the ncu launch
#                          list of the same bench command (profiles/<TAG>_* are made from these files)
#   benchn N TAG           bench.py on N GPUs launched as the driver does (gpurun --gpus N)
#   ncu    KERNEL TAG ARGS one `ncu --set full` capture of the first launch matching KERNEL in quick_bench ARGS,
#                          taken only after the same command exited 0 without ncu
#   final  TAG             last call of a round on a short budget: the newest GPU test alone, one `ncu --set full` capture of
#                          k_narrow (launches 6-8 of a C4 solve, after the same command ran without ncu), then the whole suite
#   pick   NAMES...        A/B of side builds build_timing/NAME.so against the tree's build on the C4 solve (quick_bench c4t sq), then
#                          the whole GPU suite under the fastest of them (B200LSAP_LIB) if it beats the tree's build
#   sanitize [TOOLS]       compute-sanitizer memcheck / racecheck / synccheck over tools/sanitize_cases.py
cd "${GRAFT_REPO_ROOT:-$(dirname "$0")/..}" || exit 1
mkdir -p gpurun_out
cmd=$1; shift
case "$cmd" in
quick)
  timeout 900 python tools/quick_bench.py "$@" 2>&1 | tee gpurun_out/quick.log ;;
tests)
  if [ $# -eq 0 ]; then set -- tests; fi
  timeout 1500 python -m pytest "$@" -m gpu -x -q --timeout 300 2>&1 | tee gpurun_out/pytest_gpu.log | tail -15 ;;
batch)
  { echo "== batch tests"; timeout 900 python -m pytest tests/test_gpu_batch_device.py -m gpu -x -q 2>&1 | tail -8
    echo "== c3 shards"; timeout 600 python tools/quick_bench.py c3shards "$@"; } 2>&1 | tee gpurun_out/batch.log ;;
ab)
  name=$1; shift
  { echo "== $name"; B200LSAP_LIB=$PWD/build_timing/$name.so timeout 600 python tools/quick_bench.py "$@"
    echo "== tree build"; timeout 600 python tools/quick_bench.py "$@"
    echo "== $name again"; B200LSAP_LIB=$PWD/build_timing/$name.so timeout 600 python tools/quick_bench.py "$@"; } 2>&1 | tee gpurun_out/ab_$name.log ;;
check)
  TAG=${1:-r2}
  nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt
  lscpu | grep -E "Model name|^CPU\(s\)|Thread|Core|Socket" > gpurun_out/host_cpu.txt
  timeout 1800 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log
  timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke_$TAG.log
  timeout 1200 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?" >> gpurun_out/bench_$TAG.err
  timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_$TAG.csv \
      python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/bench_under_ncu.log 2>&1
  tail -4 gpurun_out/pytest_gpu_$TAG.log; cat gpurun_out/smoke_$TAG.log; cut -c1-3000 gpurun_out/bench_$TAG.json
  tail -3 gpurun_out/bench_$TAG.err; cut -c1-1500 gpurun_out/bench_ref_$TAG.json ;;
benchn)
  N=${1:-2}; TAG=${2:-r2}
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err
  echo "rc=$?"; tail -3 gpurun_out/bench_${TAG}_n$N.err; cut -c1-3000 gpurun_out/bench_${TAG}_n$N.json ;;
ncu)
  K=$1; TAG=$2; shift 2
  timeout 600 python tools/quick_bench.py "$@" > gpurun_out/ncu_plain_$TAG.log 2>&1 && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K -s ${NCU_SKIP:-1} -c 1 -f -o gpurun_out/$TAG \
      python tools/quick_bench.py "$@" > gpurun_out/ncu_$TAG.log 2>&1
  cat gpurun_out/ncu_plain_$TAG.log; tail -3 gpurun_out/ncu_$TAG.log; ls -la gpurun_out/*.ncu-rep ;;
final)
  TAG=${1:-r2d}
  timeout 200 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q -k non_default_torch_stream > gpurun_out/pytest_new_$TAG.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/pytest_new_$TAG.log; tail -3 gpurun_out/pytest_new_$TAG.log
  timeout 120 python tools/quick_bench.py c4t > gpurun_out/ncu_plain_$TAG.log 2>&1 && \
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_narrow -s 5 -c 3 -f -o gpurun_out/${TAG}_knarrow \
      python tools/quick_bench.py c4t > gpurun_out/ncu_$TAG.log 2>&1
  cat gpurun_out/ncu_plain_$TAG.log; tail -3 gpurun_out/ncu_$TAG.log; ls -la gpurun_out/*.ncu-rep
  timeout 400 python -m pytest tests -m gpu -x -q --timeout 300 > gpurun_out/pytest_gpu_$TAG.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$TAG.log; tail -4 gpurun_out/pytest_gpu_$TAG.log ;;
pick)
  : > gpurun_out/pick.log
  best=tree; best_ms=""
  for name in tree "$@"; do
    lib=""; [ "$name" != tree ] && lib=$PWD/build_timing/$name.so
    echo "== $name" >> gpurun_out/pick.log
    B200LSAP_LIB=$lib timeout 150 python tools/quick_bench.py c4t sq >> gpurun_out/pick.log 2>&1
    ms=$(grep "dense 16384x16384" gpurun_out/pick.log | tail -1 | sed -E 's/.*: ([0-9.]+) ms rounds.*/\1/')
    echo "   -> $name c4 $ms ms" >> gpurun_out/pick.log
    case "$name" in timing*|nm*) continue ;; esac
    if [ -z "$best_ms" ]; then best_ms=$ms; tree_ms=$ms
    elif [ -n "$ms" ] && python -c "import sys; sys.exit(0 if float('$ms') < float('$best_ms') else 1)"; then best=$name; best_ms=$ms; fi
  done
  echo "== fastest: $best ($best_ms ms; tree $tree_ms ms)" >> gpurun_out/pick.log
  cat gpurun_out/pick.log
  if [ "$best" != tree ] && [ -z "$NO_SUITE" ]; then
    B200LSAP_LIB=$PWD/build_timing/$best.so timeout 400 python -m pytest tests -m gpu -x -q --timeout 300 ${PYTEST_N:+-n $PYTEST_N} > gpurun_out/pytest_gpu_pick_$best.log 2>&1
    echo "pytest rc=$?" >> gpurun_out/pytest_gpu_pick_$best.log; tail -6 gpurun_out/pytest_gpu_pick_$best.log
  fi ;;
sanitize)
  if [ $# -eq 0 ]; then set -- memcheck racecheck synccheck; fi
  for tool in "$@"; do
    timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_cases.py > gpurun_out/sanitize_$tool.log 2>&1
    echo "== $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" gpurun_out/sanitize_$tool.log | head -12; tail -2 gpurun_out/sanitize_$tool.log
  done ;;
*)
  sed -n 2,20p "$0"; exit 2 ;;
esac
