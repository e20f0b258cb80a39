for args in "256 512 20 5 2" "200 1024 40 5 1" "130 512 100 5 1" "512 1024 20 5 3 demo=1"; do
  timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) tests/mgpu_check.py $args 2>&1 | grep -E "MGPU-PARITY|differ|Error|error|Traceback|FsbError" | head -8
done
