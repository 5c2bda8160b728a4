#!/bin/bash
# Builds differently tuned copies of the library (HERE, nvcc cross-compiles) into mimas_b200/variants/.
# Usage: scripts/variants.sh name "-DMM_EXPAND_WARPS=8 ..." [name flags ...]
set -e
cd "$(dirname "$0")/../mimas_b200/csrc"
mkdir -p ../variants
while [ $# -ge 2 ]; do
  make -s OUT=../variants/libmm_$1.so EXTRA="$2" 2>&1 | grep -E "error|registers" || true
  shift 2
done
ls -la ../variants
