#!/bin/bash
# 8-GPU call of round 2: the BASELINE.json configs at their stated size and GPU count.
# C3 strong scaling at N = 2, 4, 8 runs on this one box (N = 1 comes from a 1-GPU call).
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
{ nproc; free -g; df -h /tmp /dev/shm .; nvidia-smi -L; nvidia-smi topo -m | head -12; } > $O/r2c8_box.txt 2>&1
# bit-exactness first (incl. the tests that need more than one device): a broken tree must not burn 8 GPUs' worth of minutes
if ! timeout 600 python -m pytest tests -m gpu -x -q -p no:cacheprovider \
     -k "multi or golden or mixed or short_sequences or file_backed or ordered or sharding or overflow or reupload or scaled or wide_ranges or filter" \
     > $O/r2c8_pytest.txt 2>&1; then
  echo "gpu tests failed, stopping" >> $O/r2c8_pytest.txt; tail -30 $O/r2c8_pytest.txt; exit 1
fi
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
# C3: 50 000 DNA reads, SW, fixed size at every N (BASELINE.json configs[2])
for n in 8 4 2; do
  timeout 600 $TR --nproc-per-node $n --master-port $((29510 + n)) bench.py --gpus $n --config c3 --scaling strong \
      --steps 3 --warmup 3 > $O/r2_c3_strong_n$n.json 2> $O/r2_c3_strong_n$n.err
done
# C4: 100 000 peptides, -f 0.9 (filter phase A sharded over the ranks, match lists all-gathered over NCCL), GA
timeout 900 $TR --nproc-per-node 8 --master-port 29521 bench.py --gpus 8 --config c4 --scaling strong --steps 3 --warmup 3 \
    > $O/r2_c4_n8.json 2> $O/r2_c4_n8.err
# C5: 200 000 proteins len 50..1000, SW, 80 GB packed-triangular sink in a file-backed MAP_SHARED mapping, one
# process driving all 8 devices (what the drop-in CLI's cuda_align does), strips in ascending file offset
timeout 1500 python tools/full_run.py --config c5 --ndev 8 --sink file --ordered --samples $O/r2_c5_samples.npz \
    > $O/r2_c5_full_n8.json 2> $O/r2_c5_full_n8.err
# C5-ovf: 20 000 sequences with 64 planted poly-W, pam500 (int16 overflow -> re-run) on all devices
timeout 600 python tools/full_run.py --overflow-variant --ndev 8 --sink ram --samples $O/r2_c5ovf_samples.npz \
    > $O/r2_c5ovf_n8.json 2> $O/r2_c5ovf_n8.err
# drop-in CLI with 8 visible devices on the C2 job (round 1: 5.2 s against 1.7 s with one device) and on C4 with -f
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu > $O/r2_cli_c2_8visible.json 2> $O/r2_cli_c2_8visible.err
timeout 900 python bench.py --config c4 --steps 1 --warmup 3 --no-cpu > $O/r2_cli_c4_8visible.json 2> $O/r2_cli_c4_8visible.err
ls -la $O | tail -30
