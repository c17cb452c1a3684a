#!/bin/bash
# Evidence tooling: the two graph sweeps as one launch (default) and as two (DGC_F_SPLIT_SWEEPS = 512)
for f in 0 512; do
  python bench.py --steps 3 --warmup 2 --no-cpu --cli-targets 0 --flags $f 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('flags $f value %.1f' % d['value'], d['roofline']['kernel_ms_per_step'])"
done
