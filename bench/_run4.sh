TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29614 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/bench_n4_final.json 2> gpurun_out/bench_n4_final.err; echo "ours rc=$?"; cut -c1-200 gpurun_out/bench_n4_final.json
timeout 200 $TR 29616 bench.py --gpus 4 --steps 6 --warmup 3 --comm-backend nccl > gpurun_out/bench_n4_ncclarm.json 2> gpurun_out/bench_n4_ncclarm.err; echo "nccl arm rc=$?"; cut -c1-200 gpurun_out/bench_n4_ncclarm.json
