#!/bin/bash
# round 2, final single-GPU evidence: smoke, the whole GPU test-suite, the default bench line (all workloads), the reference arm
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/final1; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread --durations=8 > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -14 $O/pytest.log
timeout 600 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "bench default rc=$?"
python -c "
import json; d=json.load(open('$O/bench_default.json')); print(round(d['value']), round(d['e2e']['value']), round(d['roofline']['frac'],3), d['roofline']['executed_flops'], d['clocks']); print({k:(round(v['value']), round(v['e2e']['value']), round(v['roofline']['frac'],4), v['detail']['lbvh']) for k,v in d.get('workloads',{}).items()})"
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"; cut -c1-300 $O/bench_reference.json
timeout 200 python bench.py --workload cornell_512_16spp_d4 --workloads none > $O/bench_c1.json 2>/dev/null; python -c "
import json; d=json.load(open('$O/bench_c1.json')); print('C1', round(d['value']), round(d['e2e']['value']), d['cpu_baseline'])"
