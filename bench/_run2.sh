TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29614 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2_final.json 2> gpurun_out/bench_n2_final.err; echo "ours rc=$?"; cut -c1-200 gpurun_out/bench_n2_final.json
timeout 200 python bench.py --gpus 1 --steps 10 --warmup 3 > gpurun_out/bench_n1_final.json 2> gpurun_out/bench_n1_final.err; echo "ours n1 rc=$?"; cut -c1-200 gpurun_out/bench_n1_final.json
