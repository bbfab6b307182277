#!/bin/bash
# GPU call 4 of round 2 (1 GPU): validation of the bit-packed subject layout (2 bits per letter for ACGT reads, 5 for
# protein) -- full GPU test suite, the default bench line, one line per other config.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c4_pytest.txt 2>&1
echo "pytest exit $?" >> $O/r2c4_pytest.txt
timeout 600 python bench.py > $O/r2c4_bench_c2.json 2> $O/r2c4_bench_c2.err
{
  tools/ab2.sh "default" --config c3 --n 20000
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter
  tools/ab2.sh "default" --config c5 --n 12000
} > $O/r2c4_ab.txt 2>&1
ls -la $O | tail -5
