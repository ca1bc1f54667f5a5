timeout 300 python bench/kernel_check.py --only attn > gpurun_out/kc_attn.log 2>&1; echo rc=$?; grep -i "fail" gpurun_out/kc_attn.log | head; grep -i "drop\|time" gpurun_out/kc_attn.log | head -20
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x 2>&1 | tail -3
timeout 200 python bench.py --gpus 1 --steps 8 --warmup 3 > gpurun_out/bench_n1_f.json 2> gpurun_out/bench_n1_f.err; echo rc=$?; cut -c1-250 gpurun_out/bench_n1_f.json
