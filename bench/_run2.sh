TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29612 bench/comm_check.py rs > gpurun_out/cc_rs.log 2>&1; echo rc=$? >> gpurun_out/cc_rs.log
grep -v "^\*\|OMP\|^$\|NCCL\|fabric" gpurun_out/cc_rs.log | tail -16
timeout 100 $TR 29613 bench/fused_trace.py > gpurun_out/ft.log 2>&1; echo rc=$? >> gpurun_out/ft.log
grep -v "^\*\|OMP\|^$\|NCCL\|fabric" gpurun_out/ft.log | tail -13
timeout 200 $TR 29614 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2_c.json 2> gpurun_out/bench_n2_c.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_n2_c.json
