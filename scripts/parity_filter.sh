#!/bin/bash
# parity pre-filter: tests + timing on/off
cd /root/repo
true
for f in 1 0; do
  PTB_PARITY_FILTER=$f python bench.py --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('filter=$f', d['value'], d.get('bit_exact_flavour'))"
done
