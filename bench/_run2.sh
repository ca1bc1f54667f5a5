TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 300 python bench.py --gpus 1 --steps 6 --warmup 3 --impl eager > gpurun_out/bench_n1_eager.json 2> gpurun_out/bench_n1_eager.err; echo "eager rc=$?"; cut -c1-250 gpurun_out/bench_n1_eager.json
timeout 300 $TR 29615 bench.py --gpus 2 --steps 6 --warmup 3 --impl eager > gpurun_out/bench_n2_eager.json 2> gpurun_out/bench_n2_eager.err; echo "eager n2 rc=$?"; cut -c1-250 gpurun_out/bench_n2_eager.json
timeout 300 $TR 29616 bench.py --gpus 2 --steps 8 --warmup 3 > gpurun_out/bench_n2_e.json 2> gpurun_out/bench_n2_e.err; echo "ours n2 rc=$?"; cut -c1-250 gpurun_out/bench_n2_e.json
