#!/usr/bin/env bash
# First GPU call of the next session: everything that was written without GPU time, in one go.
#   gpurun --timeout 1500 -- bash bench/gpu_checklist.sh            (1 GPU)
#   gpurun --gpus 2 --timeout 1500 -- bash bench/gpu_checklist.sh 2  (adds the pipeline-transport test)
# Each step runs under its own `timeout`; logs land in gpurun_out/checklist_*.log and a one-line summary per step is printed.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
NG=${1:-1}
step() { # name, seconds, command...
  local name=$1 secs=$2; shift 2
  timeout -k 5 "$secs" "$@" > "gpurun_out/checklist_${name}.log" 2>&1
  echo "[$name] exit $? : $(tail -n 1 "gpurun_out/checklist_${name}.log" | cut -c1-200)"
}
step new_kernels   900 python -m pytest tests/test_zy_conv_gpu.py tests/test_zy_softmax_gpu.py -q -rxXs
step gpu_suite    1200 python -m pytest tests -m gpu -x -q
step smoke         300 python -c "import __graft_entry__ as g; g.smoke()"
step bench_n1      600 python bench.py --steps 8 --warmup 3
step mnist_native  300 python entrypoints/pytorch_mnist.py --synthetic --epochs 1 --native-ops
step mnist_aten    300 python entrypoints/pytorch_mnist.py --synthetic --epochs 1
step gpt_variants  600 python entrypoints/pretrain_gpt.py --num-layers 12 --hidden-size 768 --num-attention-heads 12 --seq-length 1024 \
  --max-position-embeddings 1024 --micro-batch-size 12 --global-batch-size 24 --lr 3e-4 --train-iters 12 --lr-decay-iters 12 --lr-warmup-iters 2 \
  --log-interval 4 --eval-iters 0 --tokenizer-type NullTokenizer --vocab-size 32767 --mock-data --bf16 --hidden-dropout 0 --attention-dropout 0 \
  --swiglu --use-rotary-position-embeddings --group-query-attention --num-query-groups 4 --disable-bias-linear
if [ "$NG" -ge 2 ]; then
  step pp_transport 900 python -m pytest tests/test_multigpu.py -q -rxXs -k pp2
  step bench_pp     900 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$NG" --master-addr 127.0.0.1 --master-port 29511 \
    bench/bench_gpt_pp.py --gpus "$NG" --tp $((NG / 2)) --pp 2 --steps 4 --warmup 2
fi
