# one compute-sanitizer tool per call ($1 = memcheck | racecheck), smallest shapes, each section first run plain
TOOL=${1:-memcheck}
mkdir -p gpurun_out
for sec in ${2:-norm elementwise xent optim embedding gemm attn}; do
  timeout 300 python bench/kernel_check.py --quick --only $sec > gpurun_out/plain_$sec.log 2>&1 || { echo "plain $sec failed"; tail -3 gpurun_out/plain_$sec.log; continue; }
  timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 9 python bench/kernel_check.py --quick --only $sec > gpurun_out/sanitize_${TOOL}_$sec.log 2>&1
  echo "$TOOL $sec -> exit $? : $(grep 'ERROR SUMMARY' gpurun_out/sanitize_${TOOL}_$sec.log | tail -1)"
done
