#!/bin/bash
# Evidence tooling: the sweep launch on small batches (latency-bound regime) and on other shapes.
for cfg in "--targets 256" "--targets 1024" "--targets 4096" "--targets 256 --L 40000 --cov 100 --seed 5" "--targets 2400 --L 20000 --cov 50 --seed 4"; do
  python bench.py --steps 2 --warmup 2 --no-cpu --cli-targets 0 $cfg 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$cfg', 'value %.1f e2e %.1f' % (d['value'], d['e2e']['value']), {k: round(v,1) for k,v in d['roofline']['kernel_ms_per_step'].items()})"
done
