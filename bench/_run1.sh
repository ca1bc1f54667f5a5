timeout 300 python bench/kernel_check.py --only gemm > gpurun_out/kc_gemm.log 2>&1; echo rc=$?; grep -i "fail" gpurun_out/kc_gemm.log | head; grep "time\|TFLOP" gpurun_out/kc_gemm.log | head -40
timeout 200 python bench.py --gpus 1 --steps 8 --warmup 3 --dropout 0.0 > gpurun_out/bench_n1_d.json 2> gpurun_out/bench_n1_d.err; echo rc=$?; cut -c1-250 gpurun_out/bench_n1_d.json
