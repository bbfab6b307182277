#!/bin/bash
# A/B helper for the GPU box: runs bench.py with several sab_config.flags values and prints one line each.
# usage: tools/ab.sh "<flags list>" [extra bench args]
FL="$1"; shift
for f in $FL; do
  timeout 600 python bench.py --no-cpu --no-cli --steps 2 --warmup 1 --flags $f "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('flags=$f', d['config']['workload'][:70], 'value=%.0f e2e=%.0f frac=%.3f pairs/s=%.3g' % (d['value'], d['e2e']['value'], r['frac'] or 0, d['e2e']['pairs_per_s']))"
done
