#!/bin/bash
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call10; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -s --timeout 200 --timeout-method thread -k "pipelines_agree or quantised" > $O/pytest_par.log 2>&1; echo "parity rc=$?"; grep -E "fast flavour|passed|failed" $O/pytest_par.log
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q --timeout 300 --timeout-method thread > $O/pytest_multi.log 2>&1; echo "pytest multi rc=$?"; tail -3 $O/pytest_multi.log
# tile split where the ranks' shares take the pool kernel and the full frame would take the wavefront (as at 8 GPUs)
PTB_AUTO_POOL_PIXELS=5000000 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 2 --warmup 3 --workload grid1m_4k_8192spp_d4 --workloads none --spp 32 > $O/bench_tile.json 2> $O/bench_tile.err; echo "bench tile rc=$?"
python -c "
import json; d=json.load(open('$O/bench_tile.json')); print(round(d['value']), d['parity_check'])"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench 2gpu rc=$?"
python -c "
import json; d=json.load(open('$O/bench_2gpu.json')); print(round(d['value']), round(d['e2e']['value']), d['parity_check']['ok']); print({k:(round(v['value']), round(v['e2e']['value']), v['parity_check']['ok']) for k,v in d.get('workloads',{}).items()})"
