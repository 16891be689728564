#!/bin/bash
# 2 GPUs, final build: the multi-GPU tests (two processes, one handle over two devices, pipeline shards) and the headline line at N = 2
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_spmg.py tests/test_gpu_pipe.py -m gpu -q -rs > gpurun_out/pytest_multi2.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multi2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29641 \
    bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/scale_n2.json 2> gpurun_out/scale_n2.err; echo "n=2 rc=$?"
MPPI_PIPE_FOLD=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29642 \
    bench.py --gpus 2 --steps 100 --warmup 10 --no-extra > gpurun_out/scale_n2_v7.json 2> gpurun_out/scale_n2_v7.err; echo "n=2 v7 rc=$?"
python tools/summ.py gpurun_out/scale_n2.json gpurun_out/scale_n2_v7.json
python - <<'PY'
import json
for f in ('gpurun_out/scale_n2.json', 'gpurun_out/scale_n2_v7.json'):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1]); print(f, 'parity_n', d.get('parity_n'), 'cfg4', (d.get('extra') or {}).get('cfg4', {}).get('ms_per_step'))
    except Exception as e:
        print(f, 'unreadable', e)
PY
