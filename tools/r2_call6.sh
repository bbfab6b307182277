#!/bin/bash
# GPU call 6 of round 2 (1 GPU): A/B of the SAB_HOIST kernel variant (boundary exchange through shared-window
# addresses formed once per kernel, running row pointer; sab_wave.cuh) with and without row prefetch against the
# shipped build on the same box, then the parity file of the GPU suite on the variant library.
cd "$(dirname "$0")/.."
O=gpurun_out
mkdir -p $O
T0=$(date +%s)
{
  for cfg in "--config c2" "--config c3 --n 20000" "--config c4 --n 40000 --no-filter"; do
    tools/ab2.sh "default hoist hoistpf" $cfg
  done
  echo "ab done after $(( $(date +%s) - T0 )) s"
} > $O/r2c6_ab.txt 2>&1
SAB_LIB=$PWD/sequencealigner_b200/_variants/hoist.so timeout 200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -p no:cacheprovider > $O/r2c6_pytest_hoist.txt 2>&1
echo "pytest hoist exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c6_pytest_hoist.txt
if [ $(( $(date +%s) - T0 )) -lt 250 ]; then
  SAB_LIB=$PWD/sequencealigner_b200/_variants/hoistpf.so timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -p no:cacheprovider > $O/r2c6_pytest_hoistpf.txt 2>&1
  echo "pytest hoistpf exit $? after $(( $(date +%s) - T0 )) s" >> $O/r2c6_pytest_hoistpf.txt
fi
cat $O/r2c6_ab.txt; tail -2 $O/r2c6_pytest_hoist.txt
