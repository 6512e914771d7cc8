#!/bin/bash
# one 8-GPU call: pipelined pull on real ranks at N = 8 and 4, then the bench lines at N = 8, 4, 2
mkdir -p gpurun_out
for n in 8 4; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n scripts/sharded_pull_check.py > gpurun_out/sharded_pull_check_n$n.log 2>&1
  echo "pull check N=$n rc=$?"; tail -1 gpurun_out/sharded_pull_check_n$n.log | cut -c1-400
done
bash scripts/scale_run.sh "8 4 2"
