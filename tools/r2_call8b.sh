#!/bin/bash
# second 8-GPU call of round 2: what the ordered issue / the file-backed sink cost at full C5 size, and the
# multi-process lines after the strip-plan changes (equal-cost strips; geometric tail for short sequences).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
# C5 at full size, costliest-first into anonymous memory (same scores -> same checksum as the file-backed run)
timeout 900 python tools/full_run.py --config c5 --ndev 8 --sink ram > $O/r2b_c5_full_n8_ram.json 2> $O/r2b_c5_full_n8_ram.err
# ... and file-backed in ascending offsets again, now with equal-cost strips on the shared cursor
timeout 900 python tools/full_run.py --config c5 --ndev 8 --sink file --ordered > $O/r2b_c5_full_n8_file_ordered.json 2> $O/r2b_c5_full_n8_file_ordered.err
# C4 with the geometric strip tail (end-to-end is write-back bound), C2 weak scaling point (the driver's SCALE workload)
timeout 600 $TR --nproc-per-node 8 --master-port 29531 bench.py --gpus 8 --config c4 --scaling strong --steps 3 --warmup 3 \
    > $O/r2b_c4_n8.json 2> $O/r2b_c4_n8.err
timeout 600 $TR --nproc-per-node 8 --master-port 29532 bench.py --gpus 8 --steps 3 --warmup 3 > $O/r2b_c2_weak_n8.json 2> $O/r2b_c2_weak_n8.err
ls -la $O | tail -12
