#!/bin/bash
# on the GPU box: time every build/variants/*.so on the long-read workloads (kernel phases only)
for so in build/variants/*.so; do for w in c3 c4; do
  BRB_LIB=$PWD/$so timeout 400 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-workloads 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); ph=d['roofline']['phase_ms_per_step']; print('$so $w', round(d['ms_per_step'],4), [round(v,4) for v in ph.values() if v>0.001])"
done; done
