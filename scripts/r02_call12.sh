#!/bin/bash
# wavefront path state split into hot / tally / colour planes (48 B per step instead of 96): GPU tests, then the LBVH workloads
cd "$(dirname "$0")/.." || exit 1
export PYTHONUNBUFFERED=1
O=gpurun_out/call12; mkdir -p $O
timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/smoke.log
timeout 600 python -m pytest tests -m gpu -q -x --timeout 200 --timeout-method thread > $O/pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest.log
for w in spheres100k_1080p_256spp_d4 grid1m_4k_64spp_d4; do
  for pipe in wavefront; do
    timeout 300 python bench.py --workload $w --workloads none --pipeline $pipe --steps 3 --warmup 3 --no-e2e > $O/b_${w}_${pipe}.json 2>/dev/null
    python -c "
import json; d=json.load(open('$O/b_${w}_${pipe}.json')); print('$w $pipe', round(d['value']), 'parity flavour', round(d['exact']['value']) if d.get('exact') else None)"
  done
done | tee $O/sweep.txt
