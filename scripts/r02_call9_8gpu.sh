#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call9; mkdir -p $O
N=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"; tail -3 $O/bench_${N}gpu.err
python -c "
import json; d=json.load(open('$O/bench_${N}gpu.json')); print(round(d['value']), round(d['e2e']['value']), d['parity_check']['ok']); print({k:(round(v['value']), round(v['e2e']['value']), v['parity_check']['ok'], v['detail']['split']) for k,v in d.get('workloads',{}).items()})"
