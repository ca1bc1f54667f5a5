TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port"
timeout 150 $TR 29612 bench/comm_check.py agw > gpurun_out/cc4_agw.log 2>&1; echo rc=$? >> gpurun_out/cc4_agw.log
grep "FAIL\|ALL\|rc=\|graph-timed" gpurun_out/cc4_agw.log; grep -c "^ok" gpurun_out/cc4_agw.log
timeout 150 $TR 29613 bench/comm_check.py xent > gpurun_out/cc4_xent.log 2>&1; echo rc=$? >> gpurun_out/cc4_xent.log
grep "FAIL\|ALL\|rc=\|graph-timed" gpurun_out/cc4_xent.log; grep -c "^ok" gpurun_out/cc4_xent.log
