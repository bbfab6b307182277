#!/bin/bash
# GPU call 1 of round 2: box probe, full GPU test suite, kernel variant A/B, attribution runs, ncu captures.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
{ nproc; free -g; df -h /tmp /dev/shm .; python -c "import h5py; print('h5py', h5py.__version__)"; which h5dump h5ls h5cc;
  nvidia-smi -L; lscpu | grep -E "Model name|Socket|Core|Thread|MHz"; nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.limit --format=csv; } > $O/r2c1_box.txt 2>&1
timeout 1800 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c1_pytest.txt 2>&1
echo "pytest exit $?" >> $O/r2c1_pytest.txt
tools/cell_peak > $O/r2c1_cell_peak.txt 2>&1
tools/cell_peak_imm > $O/r2c1_cell_peak_imm.txt 2>&1
{ # default = per-chunk strip widths; --flags 8 / 16 = one width (32 / 16) for everything, as in round 1
  tools/ab2.sh "default imm pref pick" --config c2
  tools/ab2.sh "default" --config c2 --flags 8
  tools/ab2.sh "default imm pref" --config c3 --n 20000
  tools/ab2.sh "default" --config c3 --n 20000 --flags 8
  tools/ab2.sh "default imm pref pick" --config c4 --n 40000 --no-filter
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter --flags 16
  tools/ab2.sh "default" --config c5 --n 12000
  tools/ab2.sh "default" --config c5 --n 12000 --flags 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 1792
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 1024
  tools/ab2.sh "default" --config c2 --n 5000 --flags 1
  tools/ab2.sh "default" --config c2 --n 5000 --method nw
  tools/ab2.sh "default" --config c2 --n 5000 --method nw --flags 1
  tools/ab2.sh "default" --config c3 --n 10000 --flags 1
} > $O/r2c1_ab.txt 2>&1
timeout 900 python bench.py > $O/r2c1_bench_c2.json 2> $O/r2c1_bench_c2.err
timeout 900 python bench.py --config c4 --n 40000 --steps 2 --warmup 1 > $O/r2c1_bench_c4_40k.json 2> $O/r2c1_bench_c4_40k.err
timeout 900 python bench.py --config c3 --n 20000 --steps 2 --warmup 1 --no-cli > $O/r2c1_bench_c3_20k.json 2> $O/r2c1_bench_c3_20k.err
C4="python bench.py --config c4 --n 30000 --no-filter --no-cpu --no-cli --steps 1 --warmup 1"
$C4 > $O/r2c1_plain_c4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 6 -c 1 -o $O/r2c1_c4 $C4 > $O/r2c1_ncu_c4.log 2>&1
C3="python bench.py --config c3 --n 20000 --no-cpu --no-cli --steps 1 --warmup 1"
$C3 > $O/r2c1_plain_c3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 8 -c 1 -o $O/r2c1_c3 $C3 > $O/r2c1_ncu_c3.log 2>&1
ls -la $O
