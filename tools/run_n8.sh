N=8
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 tests/dist_parity.py --tables 16 --p2p --boundary 1000 --eval-steps 2 2>&1 | grep -E "dist parity|Error|error:|assert|rror" | tail -6 > gpurun_out/dist_parity_n8_boundary.log
cat gpurun_out/dist_parity_n8_boundary.log
