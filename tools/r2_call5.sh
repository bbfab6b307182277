#!/bin/bash
# GPU call 5 of round 2 (1 GPU): validation of the tree with the input-format parsers built into the library
# (ABI 3; the drop-in CLI now reads FASTA / DSV through sab_parse_*): full GPU test suite, smoke, the default bench
# line and the reference arm, C4 at full size through the bench incl. the CLI leg (FASTA -> filter -> align -> HDF5).
# Ordered by importance: the call may be cut short by what is left of the round's GPU budget.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
timeout 600 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c5_pytest.txt 2>&1
echo "pytest exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c5_pytest.txt
timeout 300 python bench.py > $O/r2c5_bench_c2.json 2> $O/r2c5_bench_c2.err
echo "bench exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c5_pytest.txt
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2c5_smoke.txt 2>&1
timeout 300 python bench.py --config c4 --steps 3 --warmup 3 > $O/r2c5_c4_n1.json 2> $O/r2c5_c4_n1.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2c5_bench_c2_reference.json 2> $O/r2c5_bench_c2_reference.err
echo "all done after $(( $(date +%s) - T0 )) s" >> $O/r2c5_pytest.txt
tail -3 $O/r2c5_pytest.txt; cat $O/r2c5_smoke.txt | tail -3; cut -c1-400 $O/r2c5_bench_c2.json
