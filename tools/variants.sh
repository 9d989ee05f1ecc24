#!/bin/bash
# usage: tools/variants.sh "<EXTRA flags 1>" "<EXTRA flags 2>" ...   (runs on the GPU box; rebuilds the tile kernels per
# variant and prints the headline + stage times of a 16-view slice of config C, one k_blend_jtjv launch per step)
cd "$(dirname "$0")/.."
for v in "$@"; do
  touch gaussian-splatting-lm_b200/csrc/blend_kernels.cu gaussian-splatting-lm_b200/csrc/blend_fwd.cu
  make -C gaussian-splatting-lm_b200/csrc EXTRA="$v" -j8 > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  grep -A3 "k_blend_jtjv" gaussian-splatting-lm_b200/csrc/blend_kernels.ptxas.log | grep -E "registers" | head -1
  python bench.py --steps 8 --warmup 3 --views 16 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('VARIANT [$v]', round(d['ms_per_step'],3),'ms/step (16 views)', d['stage_ms_per_step'], 'check', d['check']['norm_JtJv'])"
done
# leave the default build behind
touch gaussian-splatting-lm_b200/csrc/blend_kernels.cu gaussian-splatting-lm_b200/csrc/blend_fwd.cu
make -C gaussian-splatting-lm_b200/csrc -j8 > /dev/null 2>&1
