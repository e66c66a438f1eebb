# bench.py on N GPUs of one box, launched as the driver does (gpurun --gpus N -- bash tools/gpu_bench_n.sh N TAG)
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
N=${1:-2}; TAG=${2:-r1}
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${TAG}_n$N.json 2> gpurun_out/bench_${TAG}_n$N.err
echo "rc=$?"; tail -3 gpurun_out/bench_${TAG}_n$N.err; cut -c1-600 gpurun_out/bench_${TAG}_n$N.json
