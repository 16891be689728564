#!/bin/bash
# one box, 8 GPUs: headline config (weak scaling) and cfg4 (strong scaling, K = 1M total) at N = 8, 4 (N = 2, 1: tools/scale2.sh)
mkdir -p gpurun_out
for n in 8 4; do
  for c in 1 3; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n + c)) \
      bench.py --gpus $n --config $c --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n${n}_c${c}.log 2>&1
    echo "n=$n c=$c rc=$?"
  done
done
for n in 2; do for c in 1 3; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n + c)) \
      bench.py --gpus $n --config $c --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n${n}_c${c}.log 2>&1
done; done
for c in 1 3; do python bench.py --config $c --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n1_c${c}.log 2>&1; done
python tools/summ.py gpurun_out/scale_n*_c1.log gpurun_out/scale_n*_c3.log
