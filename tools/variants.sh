#!/bin/bash
# Tuning builds of libseqalign_b200.so: tools/variants.sh <name> [-D...]  ->  sequencealigner_b200/_variants/<name>.so
# Select at run time with SAB_LIB=<path> (sequencealigner_b200/capi.py).
set -e
cd "$(dirname "$0")/.."
python -m sequencealigner_b200.build --variant "$@"
