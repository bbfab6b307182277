#!/bin/bash
# GPU call 8 of round 2 (what is left of the budget): ncu launch list of the default bench command and one full capture
# of the dominant kernel on the final tree (both after the same commands exited 0 without ncu in call 7).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
timeout 60 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r2d_launches_c2.csv \
    python bench.py --no-cpu --no-cli --steps 2 --warmup 3 > $O/r2d_ncu_launches.log 2>&1
timeout 60 ncu --set full --clock-control none --import-source on -k regex:sab_wave -s 4 -c 1 -f -o $O/r2d_c2 \
    python bench.py --no-cpu --no-cli --steps 1 --warmup 1 > $O/r2d_ncu_c2.log 2>&1
ls -la $O | tail -5
