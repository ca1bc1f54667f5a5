TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29612 bench/comm_check.py allreduce > gpurun_out/cc_ar.log 2>&1; echo rc=$? >> gpurun_out/cc_ar.log
grep -v "^\*\|OMP\|^$\|NCCL\|fabric\|multicast" gpurun_out/cc_ar.log | tail -24
