TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port"
timeout 240 $TR 29621 bench/bench_gpt_pp.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/pp_n8_fused.json 2> gpurun_out/pp_n8_fused.err; echo "pp fused rc=$?"; cut -c1-300 gpurun_out/pp_n8_fused.json
timeout 240 $TR 29623 bench/bench_gpt_pp.py --gpus 8 --steps 5 --warmup 3 --micro-batch-size 16 > gpurun_out/pp_n8_fused_mb16.json 2> gpurun_out/pp_n8_fused_mb16.err; echo "pp fused mb16 rc=$?"; cut -c1-300 gpurun_out/pp_n8_fused_mb16.json
B2_PP_NATIVE_TP=0 timeout 240 $TR 29622 bench/bench_gpt_pp.py --gpus 8 --steps 5 --warmup 3 --micro-batch-size 16 > gpurun_out/pp_n8_nccl_mb16.json 2> gpurun_out/pp_n8_nccl_mb16.err; echo "pp nccl-tp mb16 rc=$?"; cut -c1-300 gpurun_out/pp_n8_nccl_mb16.json
timeout 200 $TR 29614 bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/bench_n8_c.json 2> gpurun_out/bench_n8_c.err; echo "bench rc=$?"; cut -c1-330 gpurun_out/bench_n8_c.json
