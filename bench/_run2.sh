TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29614 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2_g.json 2> gpurun_out/bench_n2_g.err; echo "ours rc=$?"; cut -c1-230 gpurun_out/bench_n2_g.json; grep -o '"last_loss[^}]*' gpurun_out/bench_n2_g.json
