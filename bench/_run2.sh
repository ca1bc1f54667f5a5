TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
timeout 400 python -m pytest tests/test_multigpu.py -q -x -k "zero3" 2>&1 | tail -4
timeout 300 $TR 29621 bench/bench_zero3.py --model llama-7b --layers 4 --steps 2 --warmup 2 --grad-accum 4 > gpurun_out/z3_l4_fused.json 2> gpurun_out/z3_l4_fused.err; echo rc=$?; cut -c1-200 gpurun_out/z3_l4_fused.json; grep -o '"fused_gather_gemm_units[^]]*]' gpurun_out/z3_l4_fused.json
timeout 300 $TR 29622 bench/bench_zero3.py --model llama-7b --layers 4 --steps 2 --warmup 2 --grad-accum 4 --no-fuse-gather-gemm > gpurun_out/z3_l4_plain.json 2> gpurun_out/z3_l4_plain.err; echo rc=$?; cut -c1-200 gpurun_out/z3_l4_plain.json
