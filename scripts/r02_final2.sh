#!/bin/bash
# round 2, final 2-GPU evidence: the multi-GPU tests (ptb_set_gpus, ptb_comm_*, IPC gather, the reference host on 2 GPUs), bench --gpus 2, the reference arm under torchrun
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/final2; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 300 --timeout-method thread > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -3 $O/pytest_multi.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29542 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench 2gpu rc=$?"
python -c "
import json; d=json.load(open('$O/bench_2gpu.json')); print(round(d['value']), round(d['e2e']['value']), d['parity_check']['ok']); print({k:(round(v['value']), round(v['e2e']['value']), v['parity_check']['ok']) for k,v in d.get('workloads',{}).items()})"
timeout 300 $TR --master-port 29543 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 > $O/bench_ref_2gpu.json 2> $O/bench_ref_2gpu.err; echo "bench ref under torchrun rc=$?"; cut -c1-200 $O/bench_ref_2gpu.json
