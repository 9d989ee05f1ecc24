#!/bin/bash
# usage: tools/variants.sh "<EXTRA flags 1>" "<EXTRA flags 2>" ...   (runs on the GPU box; rebuilds blend_kernels.o per variant)
cd /root/repo
for v in "$@"; do
  touch gaussian-splatting-lm_b200/csrc/blend_kernels.cu
  make -C gaussian-splatting-lm_b200/csrc EXTRA="$v" -j8 > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  grep -A2 "k_blend_jtjv" gaussian-splatting-lm_b200/csrc/blend_kernels.ptxas.log | grep -E "registers" | head -1
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('VARIANT [$v]', round(d['value'],2),'matvecs/s', d['stage_ms_per_step'])"
done
