#!/bin/bash
# usage: tools/bench_short.sh C2 C3 ...   (prints a one-line digest per config)
for c in "$@"; do python bench.py --config $c --steps 10 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']; c=d['config']
print('$c', 'value %.3e'%d['value'], 'ms/step %.4f'%d['ms_per_step'], 'score_kernel_ms %.4f'%r['kernel_ms'], 'kernel evals/s %.3e'%r['kernel_evals_per_s'], 'frac %.3f'%r['frac'], 'e2e ms %.4f'%d['e2e']['ms_per_step'], 'solved',c['hypotheses_scored_per_step'],'gb',c['guard_band_pairs'],'exact',c['exact_hypotheses'], 'cpu %.3e'%d['cpu_baseline']['value'])"; done
