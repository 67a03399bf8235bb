#!/bin/bash
# parity pre-filter: tests + timing on/off
cd /root/repo
python -m pytest tests/test_gpu_edges.py tests/test_gpu_fullsize.py tests/test_gpu_parity.py -x -q -m gpu 2>&1 | tail -3
for f in 1 0; do
  PTB_PARITY_FILTER=$f python bench.py --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('filter=$f', d['value'], d.get('bit_exact_flavour'))"
done
