N=8
: > gpurun_out/mgpu_check_n8.log
C=$((256 * N))
for args in "256 $C 20 5 2" "130 $C 100 5 1" "8192 $C 20 5 2 demo=1 halo=256" "257 $C 40 10 1" "128 $C 20 7 2 diff=2e-4"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) tests/mgpu_check.py $args 2>&1 | grep -E "MGPU-PARITY|differ|Error|error" >> gpurun_out/mgpu_check_n8.log
done
cat gpurun_out/mgpu_check_n8.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --steps 5 --warmup 3 --no-scaling-base > gpurun_out/bench_r2_n$N.json 2> gpurun_out/bench_r2_n$N.err
tail -c 200 gpurun_out/bench_r2_n$N.json; tail -2 gpurun_out/bench_r2_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 bench.py --gpus $N --steps 5 --warmup 3 --workload config4 --no-scaling-base > gpurun_out/bench_r2_c4_n$N.json 2> gpurun_out/bench_r2_c4_n$N.err
tail -c 200 gpurun_out/bench_r2_c4_n$N.json
