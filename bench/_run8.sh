TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
timeout 150 $TR 29611 bench/comm_check.py allreduce > gpurun_out/cc8_ar.log 2>&1; echo rc=$? >> gpurun_out/cc8_ar.log
grep "time\|FAIL\|ALL\|rc=" gpurun_out/cc8_ar.log
timeout 150 $TR 29612 bench/comm_check.py ar_gemm > gpurun_out/cc8_argemm.log 2>&1; echo rc=$? >> gpurun_out/cc8_argemm.log
grep "graph-timed\|FAIL\|ALL\|rc=" gpurun_out/cc8_argemm.log
timeout 150 $TR 29613 bench/bench_ddp.py --steps 10 --warmup 3 > gpurun_out/ddp_swinb_n8.json 2> gpurun_out/ddp_swinb_n8.err; echo "ddp rc=$?"; cut -c1-260 gpurun_out/ddp_swinb_n8.json
timeout 400 $TR 29614 bench/bench_zero3.py --model llama-7b --steps 2 --warmup 2 > gpurun_out/zero3_llama7b_n8_native.json 2> gpurun_out/zero3_llama7b_n8_native.err; echo "zero3 rc=$?"; cut -c1-300 gpurun_out/zero3_llama7b_n8_native.json; tail -3 gpurun_out/zero3_llama7b_n8_native.err
