#!/bin/bash
# GPU call 3 of round 2 (1 GPU, final tree): full GPU test suite, the bench lines of C2..C5 at N = 1, launch list + ncu
# captures of the dominant kernel per config, sink comparison (RAM / file-backed, ordered or not), int32 routing cost.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -p no:cacheprovider > $O/r2c3_pytest.txt 2>&1
echo "pytest exit $?" >> $O/r2c3_pytest.txt
timeout 900 python bench.py > $O/r2c3_bench_c2.json 2> $O/r2c3_bench_c2.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/r2c3_bench_c2_reference.json 2> $O/r2c3_bench_c2_reference.err
timeout 900 python bench.py --config c3 --scaling strong --steps 3 --warmup 3 --no-cli > $O/r2_c3_strong_n1.json 2> $O/r2_c3_strong_n1.err
timeout 1200 python bench.py --config c4 --steps 3 --warmup 3 > $O/r2_c4_n1.json 2> $O/r2_c4_n1.err
{ # per-GPU rate of shard 0 of an 8-GPU job next to the 1-GPU job (same kernel, same box)
  tools/ab2.sh "default" --config c2 --emulate-shards 8
  tools/ab2.sh "default" --config c2 --emulate-shards 8 --panel 1792
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter
  tools/ab2.sh "default" --config c4 --n 40000 --no-filter --flags 16
  # int32 routing / re-run cost
  tools/ab2.sh "default" --config c2 --n 5000 --flags 1
  tools/ab2.sh "default" --config c3 --n 10000 --flags 1
} > $O/r2c3_ab.txt 2>&1
# sink comparison on a 30 000-sequence C5 sample (1.8 GB packed matrix): anonymous memory, file-backed costliest-first,
# file-backed in ascending offsets
for v in "--sink ram" "--sink file" "--sink file --ordered"; do
  timeout 600 python tools/full_run.py --config c5 --n 30000 --ndev 1 $v
done > $O/r2c3_sink_c5_30k.jsonl 2> $O/r2c3_sink_c5_30k.err
timeout 600 python tools/full_run.py --overflow-variant --ndev 1 --sink ram --samples $O/r2_c5ovf_n1_samples.npz \
    > $O/r2_c5ovf_n1.json 2> $O/r2_c5ovf_n1.err
# launch list of the default bench command, then one full capture of the dominant kernel per config
B2="python bench.py --no-cpu --no-cli --steps 2 --warmup 3"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2c3_launches_c2.csv $B2 > $O/r2c3_ncu_launches.log 2>&1
C2="python bench.py --no-cpu --no-cli --steps 1 --warmup 1"
ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 4 -c 1 -o $O/r2c3_c2 $C2 > $O/r2c3_ncu_c2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 4 -c 1 -o $O/r2c3_c3 $C2 --config c3 --n 20000 > $O/r2c3_ncu_c3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 4 -c 1 -o $O/r2c3_c4 $C2 --config c4 --n 40000 --no-filter > $O/r2c3_ncu_c4.log 2>&1
ls -la $O
