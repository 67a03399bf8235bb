#!/bin/bash
# round 2, GPU call 1: full GPU test-suite, smoke, default bench line, pool-vs-wavefront on the LBVH workloads
cd "$(dirname "$0")/.." || exit 1
O=gpurun_out/call1; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_multi.py > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -30 $O/pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -8 $O/smoke.log
for pipe in wavefront pool; do
  for wl in spheres100k_1080p_256spp_d4 grid1m_4k_64spp_d4; do
    timeout 300 python bench.py --workload $wl --workloads none --pipeline $pipe --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_${wl}_${pipe}.json 2> $O/bench_${wl}_${pipe}.err
    echo "$wl $pipe rc=$?"; python -c "import json,sys; d=json.load(open('$O/bench_${wl}_${pipe}.json')); print(d['value'], d['e2e']['value'], d['bit_exact_flavour'], d['roofline']['frac'], d['roofline']['box_tests_per_ray'], d['roofline']['tri_tests_per_ray'], d['gpu_launches'])" 2>&1 | tail -1
  done
done
PTB_VERBOSE=1 timeout 300 python bench.py --workload grid1m_4k_8192spp_d4 --workloads none --pipeline pool --emulate-shard 8 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_c5_shard8_pool.json 2> $O/bench_c5_shard8_pool.err; echo "c5 shard8 pool rc=$?"
timeout 300 python bench.py --workload grid1m_4k_8192spp_d4 --workloads none --pipeline wavefront --emulate-shard 8 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > $O/bench_c5_shard8_wavefront.json 2> $O/bench_c5_shard8_wavefront.err; echo "c5 shard8 wf rc=$?"
for f in $O/bench_c5_shard8_*.json; do python -c "import json; d=json.load(open('$f')); print('$f', d['value'])"; done
timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench_default.json 2> $O/bench_default.err; echo "bench default rc=$?"
python -c "import json; d=json.load(open('$O/bench_default.json')); print(d['value'], d['e2e']['value'], d['roofline']['frac'], {k:(v['value'],v['e2e']['value']) for k,v in d.get('workloads',{}).items()})"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "bench ref rc=$?"; cat $O/bench_reference.json | cut -c1-600
