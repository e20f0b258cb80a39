# Multi-GPU round: parity cases (tests/mgpu_check.py) and the bench at N ranks.  usage: bash tools/mg_round.sh N [parity|bench|all]
N=$1
WHAT=${2:-all}
mkdir -p gpurun_out
if [ "$WHAT" != "bench" ]; then
  : > gpurun_out/mgpu_check_n$N.log
  C=$((256 * N))
  for args in "256 $C 20 5 2" "200 $((128 * N)) 40 5 1" "300 $C 33 0 2" "130 $C 100 5 1" "128 $C 20 7 2 diff=2e-4" "128 $C 10 5 2 visc=0" \
              "8192 $C 20 5 2 demo=1 halo=256" "512 $C 100 5 2" "3 $C 20 5 2" "257 $C 40 10 1" "256 $C 20 5 2 chain=0" "130 $C 100 5 1 chain=0"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) tests/mgpu_check.py $args 2>&1 | grep -E "MGPU-PARITY|differ|Error|error" >> gpurun_out/mgpu_check_n$N.log
  done
  cat gpurun_out/mgpu_check_n$N.log
fi
if [ "$WHAT" != "parity" ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err
  tail -c 300 gpurun_out/bench_r2_n$N.json; tail -3 gpurun_out/bench_r2_n$N.err
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus $N --steps 5 --warmup 3 --workload config4 --no-scaling-base > gpurun_out/bench_r2_c4_n$N.json 2> gpurun_out/bench_r2_c4_n$N.err
  tail -c 300 gpurun_out/bench_r2_c4_n$N.json
fi
