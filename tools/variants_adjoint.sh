#!/bin/bash
# usage: tools/variants_adjoint.sh "<EXTRA flags 1>" ...   (GPU box; full rebuild per variant: ACT_RANGE lives in the common header)
cd "$(dirname "$0")/.."
for v in "$@"; do
  touch gaussian-splatting-lm_b200/csrc/gslm_common.cuh
  make -C gaussian-splatting-lm_b200/csrc EXTRA="$v" -j16 > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  grep -A2 "k_preprocess_bwd_views" gaussian-splatting-lm_b200/csrc/preprocess_bwd.ptxas.log | grep -E "registers" | head -1
  python bench.py --steps 5 --warmup 3 --views 16 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('VARIANT [$v]', round(d['ms_per_step'],3),'ms/step', d['stage_ms_per_step'])"
done
touch gaussian-splatting-lm_b200/csrc/gslm_common.cuh
make -C gaussian-splatting-lm_b200/csrc -j16 > /dev/null 2>&1
