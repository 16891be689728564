#!/bin/bash
# N GPUs (default 4): headline line with the default step and with the pipeline's in-block update, alternating
N=${1:-4}
mkdir -p gpurun_out
port=29650
for r in 1 2; do for v in ${FOLDS:-0 1}; do
  port=$((port + 1))
  MPPI_PIPE_FOLD=$v timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus $N --steps 100 --warmup 10 --no-extra --no-cpu-baseline > gpurun_out/scale_n${N}_fold${v}_$r.json 2> gpurun_out/scale_n${N}_fold${v}_$r.err; echo "n=$N fold=$v rc=$?"
done; done
python tools/summ.py gpurun_out/scale_n${N}_fold*.json
