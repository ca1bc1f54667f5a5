timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -k "dropout_masks" 2>&1 | tail -5
timeout 200 python bench.py --gpus 1 --steps 8 --warmup 3 --dropout 0.1 > gpurun_out/bench_n1_drop.json 2> gpurun_out/bench_n1_drop.err; echo rc=$?; cut -c1-250 gpurun_out/bench_n1_drop.json; grep -o '"dropout[^,]*,\|"cuda_graph[^,]*,\|"last_loss[^,}]*' gpurun_out/bench_n1_drop.json
timeout 200 python bench.py --gpus 1 --steps 8 --warmup 3 --dropout 0.0 > gpurun_out/bench_n1_nodrop.json 2> gpurun_out/bench_n1_nodrop.err; echo rc=$?; cut -c1-250 gpurun_out/bench_n1_nodrop.json
