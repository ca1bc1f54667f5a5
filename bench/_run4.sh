TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29612 bench/comm_check.py allreduce > gpurun_out/cc4_ar.log 2>&1; echo rc=$? >> gpurun_out/cc4_ar.log
grep "time\|FAIL\|ALL\|rc=" gpurun_out/cc4_ar.log
