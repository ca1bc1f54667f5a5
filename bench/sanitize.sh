#!/usr/bin/env bash
# compute-sanitizer passes over the kernel tiers at small shapes.  Run on a GPU box:
#   gpurun --timeout 1500 -- bash bench/sanitize.sh            (one GPU: memcheck + racecheck of every single-GPU kernel section)
#   gpurun --gpus 2 --timeout 1500 -- bash bench/sanitize.sh 2 (adds the fabric / fused-collective checks under memcheck)
# Reports go to gpurun_out/sanitize_*.log; the script exits non-zero if any tool reports an error.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=${1:-1}
rc=0
for tool in memcheck racecheck; do
  for sec in norm elementwise xent optim embedding wattn gemm attn; do
    log=gpurun_out/sanitize_${tool}_${sec}.log
    timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 python bench/kernel_check.py --quick --only $sec > "$log" 2>&1
    code=$?
    echo "$tool $sec -> exit $code ($(grep -c 'ERROR SUMMARY: 0 errors' "$log") clean summaries)"
    [ $code -ne 0 ] && rc=1
  done
done
# kernels added late in round 1 (child-process checks)
for name in softmax_kernels conv_kernels pool_and_log_softmax batchnorm_kernels; do
  log=gpurun_out/sanitize_memcheck_${name}.log
  timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m tests._zy_checks $name > "$log" 2>&1
  code=$?
  echo "memcheck $name -> exit $code"
  [ $code -ne 0 ] && rc=1
done
if [ "$NG" -ge 2 ]; then
  log=gpurun_out/sanitize_memcheck_comm.log
  timeout 900 compute-sanitizer --tool memcheck --target-processes all --error-exitcode 9 \
    python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 --master-port 29533 bench/comm_check.py > "$log" 2>&1
  code=$?
  echo "memcheck comm_check (2 GPUs) -> exit $code"
  [ $code -ne 0 ] && rc=1
fi
exit $rc
