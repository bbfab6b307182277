#!/bin/bash
# A/B on the GPU box: tools/ab2.sh "<variant names|default>" <bench args...>; one line per variant.
V="$1"; shift
for v in $V; do
  if [ "$v" = default ]; then unset SAB_LIB; else export SAB_LIB=$PWD/sequencealigner_b200/_variants/$v.so; fi
  timeout 600 python bench.py --no-cpu --no-cli --steps 2 --warmup 1 "$@" 2>/dev/null | python tools/ab_line.py "lib=$v" "$@"
done
