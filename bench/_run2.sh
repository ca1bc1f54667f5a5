TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29611 bench/comm_check.py ag > gpurun_out/cc_ag.log 2>&1; echo rc=$? >> gpurun_out/cc_ag.log
grep "graph-timed\|FAIL\|rc=\|ALL" gpurun_out/cc_ag.log
timeout 200 $TR 29614 bench.py --gpus 2 --steps 6 --warmup 3 --dropout 0.0 > gpurun_out/bench_n2_d.json 2> gpurun_out/bench_n2_d.err; echo "bench rc=$?"; cut -c1-250 gpurun_out/bench_n2_d.json
timeout 200 $TR 29615 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2_drop.json 2> gpurun_out/bench_n2_drop.err; echo "bench dropout rc=$?"; cut -c1-250 gpurun_out/bench_n2_drop.json
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
