# Partitioned runs on N GPUs of one box: the default (component solves concurrently through the collective combiner),
# then the components in turn, each under its own timeout so that a dead-lock costs one minute, not the budget.
#   gpurun --gpus N --timeout 420 -- 'bash tools/round2_multi_gpu.sh N'
N=${1:-2}
mkdir -p gpurun_out
run() { # name, extra environment
  env $2 timeout 170 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus $N --steps 5 --warmup 3 --no-kernels --deadline 150 > gpurun_out/r2_${N}gpu_$1.json 2> gpurun_out/r2_${N}gpu_$1.err
  echo "$1 rc=$?"; tail -c 400 gpurun_out/r2_${N}gpu_$1.json
}
run default "S35_NOP=1"
[ "${2:-}" = "both" ] && run components_in_turn "S35_DIST_CONCURRENT=0"
