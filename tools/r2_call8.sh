#!/bin/bash
# 8-GPU call of round 2: the BASELINE.json configs at their stated size and GPU count.
# Everything that also runs on fewer GPUs is measured in cheaper calls.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
{ nproc; free -g; df -h /tmp /dev/shm .; nvidia-smi -L; nvidia-smi topo -m | head -12; } > $O/r2c8_box.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
# C3: 50 000 DNA reads, SW, strong scaling point N = 8 (N = 1, 2 come from 1- and 2-GPU calls)
timeout 600 $TR --master-port 29511 bench.py --gpus 8 --config c3 --scaling strong --steps 3 --warmup 3 \
    > $O/r2_c3_strong_n8.json 2> $O/r2_c3_strong_n8.err
# C4: 100 000 peptides, -f 0.9 (filter phase A sharded over the ranks), GA
timeout 600 $TR --master-port 29512 bench.py --gpus 8 --config c4 --scaling strong --steps 3 --warmup 3 \
    > $O/r2_c4_n8.json 2> $O/r2_c4_n8.err
# C5: 200 000 proteins len 50..1000, SW, 80 GB packed-triangular sink in a tmp-file mapping, one process
timeout 1200 python tools/full_run.py --config c5 --ndev 8 --sink auto --ordered --samples $O/r2_c5_samples.npz \
    > $O/r2_c5_full_n8.json 2> $O/r2_c5_full_n8.err
# drop-in CLI with 8 visible devices on the C2 job (round 1: 5.2 s against 1.7 s with one device) and on C4 with -f
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu > $O/r2_cli_c2_8visible.json 2> $O/r2_cli_c2_8visible.err
timeout 900 python bench.py --config c4 --steps 1 --warmup 1 --no-cpu > $O/r2_cli_c4_8visible.json 2> $O/r2_cli_c4_8visible.err
ls -la $O | tail -20
