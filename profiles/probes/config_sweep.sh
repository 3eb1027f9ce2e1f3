#!/bin/bash
# One bench line per BASELINE.json config on this box (1 GPU): value, e2e, cpu baseline.  Usage: bash profiles/probes/config_sweep.sh
for w in C1-pointmass C2-swimmer C3-hopper C4-ant C5-adroit C5-adroit-32; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read())
cb=d.get('cpu_baseline') or {}
print('%-14s value %12.0f ts/s  %9.3f ms/step   e2e %12.0f ts/s   cpu %10.0f ts/s (%s cores)   fvp %.3f ms' % ('$w', d['value'], d['ms_per_step'], d['e2e']['value'], cb.get('value',0), cb.get('cores','-'), d['roofline']['ms_per_launch']))"
done
