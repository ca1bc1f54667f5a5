TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29611 bench/comm_check.py ag > gpurun_out/cc4_ag.log 2>&1; echo rc=$? >> gpurun_out/cc4_ag.log
grep -v "^\*\|OMP\|^$\|NCCL\|fabric" gpurun_out/cc4_ag.log | tail -22
timeout 200 $TR 29612 bench/comm_check.py rs > gpurun_out/cc4_rs.log 2>&1; echo rc=$? >> gpurun_out/cc4_rs.log
grep -v "^\*\|OMP\|^$\|NCCL\|fabric" gpurun_out/cc4_rs.log | tail -18
timeout 200 $TR 29614 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/bench_n4_c.json 2> gpurun_out/bench_n4_c.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_n4_c.json
