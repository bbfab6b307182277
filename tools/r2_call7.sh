#!/bin/bash
# GPU call 7 of round 2 (1 GPU, last of the round): the tree with the hoisted boundary exchange as the NW / GA build --
# full GPU test suite, smoke, the default bench line, one line each for C3 (SW: unchanged images) and C4.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
timeout 240 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c7_pytest.txt 2>&1
echo "pytest exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c7_pytest.txt
timeout 120 python bench.py > $O/r2c7_bench_c2.json 2> $O/r2c7_bench_c2.err
echo "bench exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c7_pytest.txt
{
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter
  tools/ab2.sh "default" --config c3 --n 20000
  tools/ab2.sh "default" --config c2 --method nw
} > $O/r2c7_ab.txt 2>&1
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2c7_smoke.txt 2>&1
echo "all done after $(( $(date +%s) - T0 )) s" >> $O/r2c7_pytest.txt
tail -4 $O/r2c7_pytest.txt; cat $O/r2c7_ab.txt; tail -3 $O/r2c7_smoke.txt; cut -c1-300 $O/r2c7_bench_c2.json
