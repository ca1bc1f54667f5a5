timeout 300 python bench/kernel_check.py --only attn > gpurun_out/kc_attn.log 2>&1; echo rc=$?; grep -c "^ok" gpurun_out/kc_attn.log; grep "^FAIL\|^time" gpurun_out/kc_attn.log | head
timeout 600 python -m pytest tests/test_multigpu.py -q -x -k "distributed_optimizer or zero3" 2>&1 | tail -5
