#!/bin/bash
# GPU call 2 of round 2: specialised device images (immediate gap constant), concurrent per-width launches,
# SW row prefetch, equal-cost row strips (auto panels; --panel N = uniform as before: 1792 / 640 were the old auto
# choices for C2 at 1 / 8 shards).  Full GPU test suite, A/B lines, counters of the bare DP block.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c2_pytest.txt 2>&1
echo "pytest exit $?" >> $O/r2c2_pytest.txt
{ # default = per-chunk widths on concurrent side streams; 512 = the same launches back to back; 8 / 16 = one width
  tools/ab2.sh "default" --config c2
  tools/ab2.sh "default" --config c2 --flags 512
  tools/ab2.sh "default" --config c2 --flags 8
  tools/ab2.sh "default nopref prefall" --config c3 --n 20000
  tools/ab2.sh "default" --config c3 --n 20000 --flags 8
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter --flags 512
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter --flags 16
  tools/ab2.sh "default prefall" --config c5 --n 12000
  tools/ab2.sh "default" --config c5 --n 12000 --flags 512
  tools/ab2.sh "default" --config c5 --n 12000 --flags 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --flags 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 1792
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 1792 --flags 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 3584 --flags 8
  tools/ab2.sh "default" --config c2 --panel 1792
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 640
  tools/ab2.sh "default" --config c2 --panel 2560
  tools/ab2.sh "default" --config c2 --panel 2560 --flags 8
  tools/ab2.sh "default" --config c2 --panel 3328 --flags 8
} > $O/r2c2_ab.txt 2>&1
timeout 900 python bench.py > $O/r2c2_bench_c2.json 2> $O/r2c2_bench_c2.err
# counters behind "3 DPX + 1 IMAD + 1 LDS costs 7.5 clk, not 6": the bare 8 x 32 block, register vs immediate gap constant
for v in cell_peak cell_peak_imm; do
  ncu --set full --clock-control none --import-source on -k regex:micro -c 1 -o $O/r2c2_$v tools/$v > $O/r2c2_ncu_$v.log 2>&1
done
C2="python bench.py --no-cpu --no-cli --steps 1 --warmup 1"
$C2 > $O/r2c2_plain_c2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 6 -c 1 -o $O/r2c2_c2 $C2 > $O/r2c2_ncu_c2.log 2>&1
ls -la $O
