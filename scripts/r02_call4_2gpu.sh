#!/bin/bash
# 2 GPUs: the multi-GPU tests (one process per GPU over NCCL + CUDA IPC; ptb_set_gpus in one process; the reference host on 2 GPUs), bench --gpus 2
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call4; mkdir -p $O
nvidia-smi -L > $O/smi.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 300 --timeout-method thread > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -40 $O/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench 2gpu rc=$?"; tail -5 $O/bench_2gpu.err
python -c "
import json; d=json.load(open('$O/bench_2gpu.json')); print(d['value'], d['e2e']['value'], d['parity_check']); print({k:(round(v['value']), round(v['e2e']['value']), v['parity_check']) for k,v in d.get('workloads',{}).items()})"
