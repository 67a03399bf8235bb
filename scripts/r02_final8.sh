#!/bin/bash
# round 2, final multi-GPU evidence: the default bench line at N GPUs (strong scaling, all workloads) and ONE full C5 step (8192 spp)
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
N=${1:-8}
O=gpurun_out/final$N; mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
timeout 500 $TR bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_${N}gpu.json 2> $O/bench_${N}gpu.err; echo "bench ${N}gpu rc=$?"; tail -3 $O/bench_${N}gpu.err
python -c "
import json; d=json.load(open('$O/bench_${N}gpu.json')); print(round(d['value']), round(d['e2e']['value']), d['parity_check']['ok']); print({k:(round(v['value']), round(v['e2e']['value']), v['parity_check']['ok'], v['detail']['split']) for k,v in d.get('workloads',{}).items()})"
if [ "${2:-full}" = full ]; then
timeout 400 $TR bench.py --gpus $N --workload grid1m_4k_8192spp_d4 --workloads none --spp 8192 --steps 1 --warmup 1 > $O/bench_c5_full_${N}gpu.json 2> $O/bench_c5_full_${N}gpu.err; echo "C5 full rc=$?"; tail -3 $O/bench_c5_full_${N}gpu.err
python -c "
import json; d=json.load(open('$O/bench_c5_full_${N}gpu.json')); print('C5 full', round(d['value']), round(d['e2e']['value']), d['ms_per_step'], d['parity_check'], d['config'])"
fi
