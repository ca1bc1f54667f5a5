TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29612 bench/comm_check.py xent > gpurun_out/cc_xent.log 2>&1; echo rc=$? >> gpurun_out/cc_xent.log
grep "graph-timed\|FAIL\|ALL\|rc=" gpurun_out/cc_xent.log
