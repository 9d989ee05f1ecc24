cd /root/repo
for v in orig new; do
  cp tools/ab/blend_sweeps_$v.cuh gaussian-splatting-lm_b200/csrc/blend_sweeps.cuh
  make -C gaussian-splatting-lm_b200/csrc -j8 > /dev/null 2>&1 || echo build failed
  for views in 16 0; do
  python bench.py --steps 6 --warmup 3 --views $views --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$v views=$views', round(d['ms_per_step'],3), d['stage_ms_per_step'])"
  done
done
