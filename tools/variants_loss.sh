#!/bin/bash
# usage: tools/variants_loss.sh "<EXTRA flags 1>" "<EXTRA flags 2>" ...   (runs on the GPU box; rebuilds loss_blocks.o per variant)
cd "$(dirname "$0")/.."
for v in "$@"; do
  touch gaussian-splatting-lm_b200/csrc/loss_blocks.cu
  make -C gaussian-splatting-lm_b200/csrc EXTRA="$v" -j8 > /dev/null 2>&1 || { echo "build failed: $v"; continue; }
  grep "registers" gaussian-splatting-lm_b200/csrc/loss_blocks.ptxas.log | tr '\n' ' '; echo
  python tools/bench_loss.py --no-torch 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('VARIANT [$v]', {k: round(v,3) for k,v in d.items() if k.startswith('ours')})"
done
touch gaussian-splatting-lm_b200/csrc/loss_blocks.cu
make -C gaussian-splatting-lm_b200/csrc -j8 > /dev/null 2>&1
