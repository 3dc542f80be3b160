#!/bin/bash
# Everything that needs N > 1 GPUs of one box, in one go (N = first argument, default: all visible):
#   1. the real multi-GPU pytest cases (IPC / NVLS / hooks / launcher / missing-peer), log kept
#   2. bench.py --gpus N: this arm, parity block, stock-GPU leg, loss parity
#   3. the launch-shape sweep of the zero-copy all-reduce kernels
# Outputs land in gpurun_out/ with an r2_w<N>_ prefix; copy what is to be kept into profiles/.
set -u
N=${1:-$(nvidia-smi -L | wc -l)}
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
(timeout 600 python -m pytest tests/test_gpu_allreduce.py -q -k "real_peer or launcher or missing_peer" -v > $OUT/r2_w${N}_pytest.log 2>&1; echo "pytest rc=$?" >> $OUT/r2_w${N}_pytest.log)
tail -4 $OUT/r2_w${N}_pytest.log
timeout 900 $TR --master-port 29521 bench.py --gpus $N --steps ${STEPS:-10} --warmup 3 > $OUT/r2_w${N}_bench.json 2> $OUT/r2_w${N}_bench.err; echo "bench rc=$?"
tail -c 400 $OUT/r2_w${N}_bench.err
timeout 400 $TR --master-port 29522 benchmarks/microbench.py allreduce_tune --sizes-mb 8 25 64 128 --iters 20 > $OUT/r2_w${N}_tune.jsonl 2> $OUT/r2_w${N}_tune.err; echo "tune rc=$?"
tail -c 300 $OUT/r2_w${N}_tune.err
