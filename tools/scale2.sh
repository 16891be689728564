#!/bin/bash
# 2 GPUs: multi-GPU parity test, then headline config (weak) and cfg4 (strong) at N = 2 and 1
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -n 3
for c in 1 3; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29610 + c)) \
    bench.py --gpus 2 --config $c --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n2_c${c}.log 2>&1
  echo "n=2 c=$c rc=$?"
  python bench.py --config $c --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/scale_n1_c${c}.log 2>&1
done
python tools/summ.py gpurun_out/scale_n1_c1.log gpurun_out/scale_n2_c1.log gpurun_out/scale_n1_c3.log gpurun_out/scale_n2_c3.log
