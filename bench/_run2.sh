TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29614 bench.py --gpus 2 --steps 6 --warmup 3 --kernel-profile gpurun_out/kprof_n2_b.txt > gpurun_out/bench_n2_b.json 2> gpurun_out/bench_n2_b.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_n2_b.json
head -12 gpurun_out/kprof_n2_b.txt | cut -c1-200
timeout 600 python -m pytest tests/test_multigpu.py -q -x 2>&1 | tail -5
