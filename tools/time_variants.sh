#!/bin/bash
# on the GPU box: time every build/variants/*.so with the default bench workload (kernel phases only; the bench-size oracle check runs too)
for so in build/variants/*.so; do
  BRB_LIB=$PWD/$so timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e --no-workloads 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); ph=d['roofline']['phase_ms_per_step']; print('$so', round(d['ms_per_step'],4), [round(v,4) for v in ph.values() if v>0.01])"
done
