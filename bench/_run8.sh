TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
timeout 200 $TR 29614 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/bench_n8_final.json 2> gpurun_out/bench_n8_final.err; echo "ours rc=$?"; cut -c1-200 gpurun_out/bench_n8_final.json
timeout 200 $TR 29615 bench.py --gpus 8 --steps 10 --warmup 3 --dropout 0 > gpurun_out/bench_n8_nodrop.json 2> gpurun_out/bench_n8_nodrop.err; echo "ours no-dropout rc=$?"; cut -c1-200 gpurun_out/bench_n8_nodrop.json
timeout 200 $TR 29616 bench.py --gpus 8 --steps 6 --warmup 3 --comm-backend nccl > gpurun_out/bench_n8_ncclarm.json 2> gpurun_out/bench_n8_ncclarm.err; echo "nccl arm rc=$?"; cut -c1-200 gpurun_out/bench_n8_ncclarm.json
timeout 200 $TR 29617 bench.py --gpus 8 --steps 6 --warmup 3 --impl eager > gpurun_out/bench_n8_eager.json 2> gpurun_out/bench_n8_eager.err; echo "eager rc=$?"; cut -c1-200 gpurun_out/bench_n8_eager.json
