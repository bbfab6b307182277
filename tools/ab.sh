#!/bin/bash
# A/B helper for the GPU box: runs bench.py with several sab_config.flags values and prints one line each.
# usage: tools/ab.sh "<flags list>" [extra bench args]
FL="$1"; shift
for f in $FL; do
  timeout 300 python bench.py --no-cpu --steps 2 --warmup 1 --flags $f "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('flags=$f', d['config']['workload'][:60], 'value=%.0f e2e=%.0f launches=%d' % (d['value'], d['e2e']['value'], d['gpu_launches']))"
done
